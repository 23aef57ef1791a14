"""Developer aid: summarise an `ncu --page source --csv` export (SASS view) -- top instructions by stall samples and
the dominant stall reason of each.  usage: dev_ncu_source.py src.csv [lo hi]  (print SASS rows lo..hi)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
body = [r for r in rows[2:] if len(r) == len(hdr)]
ci = {h: i for i, h in enumerate(hdr)}
S = ci["# Samples"]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
def f(x):
    try: return float(x)
    except ValueError: return 0.0
tot = sum(f(r[S]) for r in body)
print("instructions", len(body), "samples", tot)
if len(sys.argv) > 3:
    lo, hi = int(sys.argv[2]), int(sys.argv[3])
    for k in range(lo, hi):
        r = body[k]
        top = max(stalls, key=lambda h: f(r[ci[h]]))
        print(k, r[ci["Source"]][:90].ljust(90), int(f(r[S])), int(f(r[ci["Instructions Executed"]])), top if f(r[S]) else "")
else:
    order = sorted(range(len(body)), key=lambda k: -f(body[k][S]))[:60]
    for k in sorted(order):
        r = body[k]
        top = sorted(stalls, key=lambda h: -f(r[ci[h]]))[:2]
        print(k, r[ci["Source"]][:80].ljust(80), int(f(r[S])), "%.1f%%" % (100 * f(r[S]) / tot), [(h, int(f(r[ci[h]]))) for h in top])
