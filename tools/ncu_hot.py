"""top SASS instructions by stall samples from `ncu --page source --csv` output
usage: ncu -i rep --page source --csv --kernel-name regex:NAME | python tools/ncu_hot.py [N]"""
import csv, sys, collections
rows = list(csv.reader(sys.stdin))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 25
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
col = {h: i for i, h in enumerate(hdr)}
S = col["# Samples"]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = collections.Counter()
total = 0
for r in body:
    total += int(r[S] or 0)
    for h in stalls:
        tot[h] += int(r[col[h]] or 0)
print("total samples", total, "instructions", len(body))
print("stall totals:", ", ".join("%s=%.1f%%" % (k[6:], 100.0 * v / max(1, total)) for k, v in tot.most_common(8)))
ex = col["Instructions Executed"]
print("inst executed total", sum(int(r[ex] or 0) for r in body))
order = sorted(range(len(body)), key=lambda i: -int(body[i][S] or 0))[:n]
for i in sorted(order):
    r = body[i]
    top = sorted(((int(r[col[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
    print("%5d %6s smp %8s exec  confl %-4s %-60s %s" % (i, r[S], r[ex], r[col["L1 Conflicts Shared N-Way"]], r[1].strip()[:60], top))
