"""CPU replay of the phase-B schedule of k_win_linearize_schur on one C5 window (no GPU): how many of the lane-rounds
the 8-lane sub-warps spend on pairs, for the block order as built, for blocks sorted by pair count, and with the last
pass split in halves.  A tile is cut like ba_setup.cu cuts it (<= 88 points, <= 640 observations);
16 sub-warps per CTA, four per warp; a warp's round count in a pass is the largest of its four sub-warps' lists."""
import math
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import viso_b200 as vb  # noqa: E402  (host-only synthetic generator)

p = vb.synth_config(5, 0)
ep, el = p["edge_pose"], p["edge_point"]
Np, nc = len(p["points_init"]), len(p["poses_init"])
vis = np.zeros((Np, nc), bool)
vis[el, ep] = True
nf = nc - 1  # pose 0 is the gauge
cnt = vis.sum(1)
tiles, p0 = [], 0
while p0 < Np:
    n = o = 0
    while p0 + n < Np and n < 88 and o + cnt[p0 + n] <= 640:
        o += cnt[p0 + n]
        n += 1
    tiles.append((p0, n))
    p0 += n


def sched(counts, nsub=16):
    rounds = 0
    for ps in range(0, len(counts), nsub):
        chunk = list(counts[ps:ps + nsub]) + [0] * (nsub - len(counts[ps:ps + nsub]))
        for w in range(4):
            rounds += max(math.ceil(c / 8) for c in chunk[4 * w:4 * w + 4])
    return rounds


useful = cur = srt = split = 0.0
for p0, n in tiles:
    v = vis[p0:p0 + n, 1:].astype(int)
    C = v.T @ v
    blocks = [C[i, i] for i in range(nf)] + [C[i, j] for i in range(nf) for j in range(i)]
    useful += sum(blocks) / 32.0
    cur += sched(blocks)
    sb = sorted(blocks, reverse=True)
    srt += sched(sb)
    k = (len(sb) // 16) * 16
    halves = [h for c in sb[k:] for h in ((c + 1) // 2, c // 2)]
    split += sched(sb[:k]) + sched(halves)
print("window: %d poses, %d points, %d observations, %d tiles, %d S blocks" % (nc, Np, len(ep), len(tiles), nf * (nf + 1) // 2))
print("useful lane-rounds / scheduled:  as built %.3f   sorted by pair count %.3f   sorted + last pass split %.3f"
      % (useful / cur, useful / srt, useful / split))
