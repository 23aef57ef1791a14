"""Summarise the ncu outputs of scripts/ncu_round2.sh into the small CSVs committed under profiles/ (run here, no GPU):
   python scripts/ncu_summarize.py <tag>   reads gpurun_out/<tag>_launches.csv, <tag>_dec.ncu-rep, <tag>_conv.ncu-rep"""
import collections, csv, json, subprocess, sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r2e"
G, P = "gpurun_out/", "profiles/"

rows = [r for r in csv.reader(open(f"{G}{tag}_launches.csv")) if len(r) > 10]
h = rows[0]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v = v / 1e6 if r[ui] == "ns" else v / 1e3 if r[ui] == "us" else v
    agg[r[ki]][0] += 1
    agg[r[ki]][1] += v
tot = sum(v for _, v in agg.values())
with open(f"{P}launches_{tag}_summary.csv", "w") as f:
    f.write(f"# ncu launch list summary ({tag}): bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-extras; "
            "gpu__time_duration.sum, --clock-control none\n# per-launch times are cold-cache and serialised: compare SHARES, not absolutes\n"
            "kernel,launches,total_ms,share\n")
    for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"\"{k[:90]}\",{n},{v:.3f},{v / tot:.4f}\n")


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(out.splitlines()))


KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "launch__cluster_size")
r = raw(f"{G}{tag}_dec.ncu-rep")
vals = dict(zip(r[0], zip(r[1], r[2])))
with open(f"{P}ncu_full_{tag}_decoder_persistent_kernel.csv", "w") as f:
    f.write("# ncu --set full --clock-control none --import-source on -k regex:decoder_persistent -s 3 -c 1, bench.py workload "
            "(B=256, T=200, 500 steps)\nmetric,unit,value\n")
    for k in KEEP:
        if k in vals:
            f.write(f"{k},{vals[k][0]},{vals[k][1]}\n")
rd, wr = float(vals["dram__bytes_read.sum"][1]), float(vals["dram__bytes_write.sum"][1])
mult = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
rd, wr = rd * mult[vals["dram__bytes_read.sum"][0]], wr * mult[vals["dram__bytes_write.sum"][0]]
json.dump({"decoder_persistent_kernel": {"dram_bytes_per_launch": rd + wr, "dram_read_bytes": rd, "dram_write_bytes": wr,
                                         "decoder_steps_per_launch": 500,
                                         "source": f"ncu --set full --clock-control none, profiles/ncu_full_{tag}_decoder_persistent_kernel.csv "
                                                   "(bench.py workload, B=256, T=200)"}},
          open(f"{P}ncu_traffic.json", "w"), indent=1)
src = subprocess.run(["ncu", "-i", f"{G}{tag}_dec.ncu-rep", "--page", "source", "--csv"], capture_output=True, text=True).stdout
rs = list(csv.reader(src.splitlines()))
hd = rs[1]
body = [x for x in rs[2:] if len(x) == len(hd)]
ci = {n: i for i, n in enumerate(hd)}
st = [n for n in hd if n.startswith("stall_") and "Not Issued" not in n]
fl = lambda x: float(x) if x.replace(".", "", 1).isdigit() else 0.0
totals = {n: sum(fl(x[ci[n]]) for x in body) for n in st}
T = sum(totals.values())
inst = sum(fl(x[ci["Instructions Executed"]]) for x in body)
with open(f"{P}ncu_source_{tag}_decoder_stall_shares.csv", "w") as f:
    f.write(f"# warp-state samples of decoder_persistent_kernel by stall reason ({tag}; --page source of the same capture)\n"
            f"# total samples {int(T)}; warp instructions executed {int(inst)} ({inst / 148 / 500:.0f} per SM per decoder step)\n"
            "stall_reason,samples,share\n")
    for n, v in sorted(totals.items(), key=lambda kv: -kv[1]):
        f.write(f"{n},{int(v)},{v / T:.4f}\n")
r = raw(f"{G}{tag}_conv.ncu-rep")
with open(f"{P}ncu_full_{tag}_conv_tc_kernel.csv", "w") as f:
    f.write("# ncu --set full --clock-control none -k regex:conv_tc_kernel -s 9 -c 3, bench.py workload (B=256, T=200): the launch "
            "with grid 6912 (= 16 widths x 432 row tiles) is the one-launch conv bank\nlaunch,metric,unit,value\n")
    for li, vv in enumerate(r[2:5]):
        d = dict(zip(r[0], zip(r[1], vv)))
        for k in KEEP:
            if k in d:
                f.write(f"{li},{k},{d[k][0]},{d[k][1]}\n")
print(open(f"{P}launches_{tag}_summary.csv").read()[:700])
print(open(f"{P}ncu_source_{tag}_decoder_stall_shares.csv").read()[:500])
