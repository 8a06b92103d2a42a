"""split an `ncu --page source --csv` dump into runs of equal execution count (loop bodies) and
print each run's share of instructions, stall samples and shared-memory wavefronts"""
import csv, sys
rows = list(csv.reader(sys.stdin))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
col = {h: i for i, h in enumerate(hdr)}
X, S, W = col["Instructions Executed"], col["# Samples"], col["L1 Wavefronts Shared"]
iv = lambda r, c: int(r[c] or 0)
tx, ts, tw = (sum(iv(r, c) for r in body) for c in (X, S, W))
minlen = int(sys.argv[1]) if len(sys.argv) > 1 else 30
runs = []; start = 0
for i in range(1, len(body) + 1):
    if i == len(body) or abs(iv(body[i], X) - iv(body[start], X)) > 0.25 * max(1, iv(body[start], X)):
        runs.append((start, i)); start = i
# merge short runs into neighbours for readability
merged = []
for a, b in runs:
    if merged and (b - a) < minlen:
        merged[-1] = (merged[-1][0], b)
    else:
        merged.append((a, b))
for a, b in merged:
    ex = sum(iv(r, X) for r in body[a:b]); sm = sum(iv(r, S) for r in body[a:b]); wv = sum(iv(r, W) for r in body[a:b])
    print("lines %5d-%5d  exec/line %9d  inst %5.1f%%  samples %5.1f%%  smem %5.1f%%" % (a, b, iv(body[a], X), 100 * ex / tx, 100 * sm / ts, 100 * wv / max(1, tw)))
