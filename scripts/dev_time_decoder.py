"""Developer aid: time the persistent decoder kernel alone (CUDA events) and print the
per-phase cycle profile.  Not part of the bench contract."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import sat_b200
from sat_b200.engine import DecoderHandle

def run(cfg, B, T, S, feed, reps=3):
    hp = sat_b200.config(cfg)
    W = sat_b200.init_weights(hp, seed=12345)
    rng = np.random.default_rng(0)
    dims = [hp.cbhg_out_units] if hp.decoder == "ExtendedDecoder" else [hp.cbhg_out_units, hp.self_attention_out_units]
    dev = torch.device("cuda:0")
    mems = [torch.from_numpy(rng.normal(0, .5, size=(B, T, d)).astype(np.float32)).to(dev) for d in dims]
    lengths = torch.from_numpy(rng.integers(T * 2 // 3, T + 1, size=B).astype(np.int64)).to(dev)
    target = torch.from_numpy(rng.normal(0, 1, size=(B, S * hp.outputs_per_step, hp.num_mels)).astype(np.float32)).to(dev)
    h = DecoderHandle(hp, W)
    kw = dict(target=target, feed_mode=feed, stop_mode=int(os.environ.get("SAT_DEV_STOP", "0")), max_steps=S)
    o = h.run(mems, lengths, **kw); torch.cuda.synchronize()
    print("n_steps", int(o["n_steps"].item()), "rows stopped", int((o["stop_steps"] >= 0).sum().item()))
    h.profile(reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        h.run(mems, lengths, **kw)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    prof = h.profile()
    tot = sum(prof) or 1
    by = h.step_bytes(B, T, S // 2)
    print(f"{cfg} B={B} T={T} S={S} feed={feed}: {ms:.2f} ms/run, {ms*1e3/S:.1f} us/step, "
          f"{B*S*hp.outputs_per_step/ms*1e3:.3e} frames/s, bytes/step {by/1e6:.1f} MB -> {by/(ms*1e-3/S)/1e9:.0f} GB/s")
    tot = sum(prof[:16]) or 1
    clk = sum(prof[:16]) / (ms * 1e3 * reps)  # SM cycles per us, from CTA 0's phase counters over the event time
    print(f"  SM clock ~{clk:.0f} MHz")
    print("  phase us/step:", " ".join(f"{p/clk/S/reps:.1f}" for p in prof[:16] if p))
    print("  attention stages us/step (CTA0):", " ".join(f"{p/clk/S/reps:.1f}" for p in prof[16:24]))
    print("  frame stages us/step (CTA0): load y, out proj, stores+feed, prenet0, prenet1(+ms), store:", " ".join(f"{p/clk/S/reps:.1f}" for p in prof[32:38]))
    if prof[30]:
        nb = int(os.environ.get("SAT_BENCH_BARRIERS", "0"))
        print("  empty grid barrier: %.2f us each" % (prof[30] / clk / nb / reps))
    print("  tc lstm (3 phases summed) us/step: producer %.1f, until accumulator ready %.1f, cell update %.1f" % tuple(p/clk/S/reps for p in prof[24:27]))

if __name__ == "__main__" and len(sys.argv) > 1:
    run(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), reps=1)
elif __name__ == "__main__":
    run("ljspeech_tacotron", 256, 200, 500, 1)
    run("ljspeech_tacotron", 16, 150, 400, 0)
    run("ljspeech_self_attention", 16, 150, 400, 0)
    run("ljspeech_self_attention", 128, 150, 400, 0)
    run("ljspeech_self_attention", 256, 200, 500, 1)
