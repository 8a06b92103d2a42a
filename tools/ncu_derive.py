#!/usr/bin/env python
"""Derive the roofline constants bench.py reports from an `ncu --set full` capture of the CURRENT build
(run here, on the CPU box, after gpurun brought the report back):

    python tools/ncu_derive.py gpurun_out/prof_win.ncu-rep --problems 1024 --observations 15377003 \
        --capture profiles/r2a_win_c5_full.csv

writes profiles/derived_roofline.json:
    {"source_digest": <viso_b200/.libviso_ba.stamp>, "capture": ..., "problems": N, "observations": No,
     "kernels": {"win_linearize_schur": {"dram_bytes": read + write per launch, "duration_us": ...,
                 "fp64_inst_per_observation": warp-level FP64-pipe instructions x 32 / No, "registers": ...}, ...}}

Per kernel the LARGEST captured launch is taken (the full-grid one: straggler launches are tiny).  bench.py reads
the file for `roofline.traffic` and `roofline.fp64.slots_per_observation`, and reports `capture_is_of_this_build`
= (digest of the library it runs == source_digest)."""
import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = {"k_win_linearize_schur": "win_linearize_schur", "k_win_finalize": "win_finalize", "k_win_update": "win_update",
         "k_solve_small": "solve_small", "k_linearize": "linearize", "k_schur_blocks": "schur_blocks",
         "k_block_finalize": "block_finalize", "k_update": "update", "k_win_single": "win_single"}


def num(x):
    try:
        return float(x.replace(",", ""))
    except Exception:
        return None


def scale(unit):
    u = unit.lower()
    return {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3,
            "msecond": 1e3, "nsecond": 1e-3, "second": 1e6, "s": 1e6}.get(u, 1.0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--problems", type=int, required=True)
    ap.add_argument("--observations", type=int, required=True, help="observations one full launch processes")
    ap.add_argument("--capture", default=None, help="name of the committed summary of the same report (profiles/...)")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "derived_roofline.json"))
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}

    def get(r, name):
        i = col.get(name)
        if i is None:
            return None
        v = num(r[i])
        return None if v is None else v * scale(units[i])

    best = {}
    for r in rows[2:]:
        k = r[col["Kernel Name"]].split("(")[0]
        name = NAMES.get(k)
        if not name:
            continue
        dur = get(r, "gpu__time_duration.sum")
        if name in best and best[name]["duration_us"] >= (dur or 0):
            continue
        rd, wr = get(r, "dram__bytes_read.sum"), get(r, "dram__bytes_write.sum")
        fp64 = get(r, "sm__inst_executed_pipe_fp64.sum")
        if fp64 is None:
            fp64 = get(r, "smsp__inst_executed_pipe_fp64.sum")
        how = "sm__inst_executed_pipe_fp64.sum (warp instructions)"
        if fp64 is None:  # --set full alone has the pipe's busy fraction only: one FP64 warp instruction holds an SMSP's pipe 2 cycles
            pct, cyc = get(r, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"), get(r, "smsp__cycles_active.avg")
            if pct is not None and cyc is not None:
                fp64 = pct / 100.0 * cyc * 148 * 4 / 2.0
                how = "sm__pipe_fp64_cycles_active% x smsp__cycles_active x 592 SMSPs / 2 cycles per warp instruction"
        rec = {"duration_us": dur, "dram_bytes": (rd or 0) + (wr or 0), "grid": r[col["Grid Size"]],
               "registers": get(r, "launch__registers_per_thread"),
               "fp64_pipe_active_pct": get(r, "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"),
               "warps_active_pct": get(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
               "fp64_warp_inst": fp64, "fp64_inst_source": how,
               "fp64_inst_per_observation": (fp64 * 32.0 / a.observations) if fp64 else None}
        best[name] = rec
    try:
        with open(os.path.join(ROOT, "viso_b200", ".libviso_ba.stamp")) as fh:
            digest = fh.read().strip()
    except Exception:
        digest = None
    out = {"source_digest": digest, "capture": a.capture or os.path.basename(a.report), "problems": a.problems,
           "observations": a.observations, "kernels": best,
           "_note": "written by tools/ncu_derive.py from one ncu --set full capture of the default bench command; "
                    "cold-cache, serialised launches: use the byte and instruction COUNTS, not the durations"}
    with open(a.out, "w") as fh:
        json.dump(out, fh, indent=1)
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
