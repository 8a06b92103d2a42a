#!/usr/bin/env python
"""SASS evidence for profiles/ (runs here, no GPU): per kernel of libviso_ba.so the opcode histogram (DFMA / DMUL /
DADD, LDS widths, LDGSTS, UBLKCP, shuffles, barriers), local-memory (spill) instructions, and the resource usage
`cuobjdump --dump-resource-usage` reports (registers, shared memory, stack).  Full listings of the hot kernels go to
profiles/sass/<kernel>.sass, the summary to profiles/<tag>_sass_summary.json.

    python tools/sass_summary.py r2a k_win_linearize_schur k_win_update k_chol_coop ...
"""
import collections
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "viso_b200", "libviso_ba.so")
INTEREST = ("DFMA", "DMUL", "DADD", "DSETP", "MUFU", "LDS", "STS", "LDG", "STG", "LDGSTS", "UBLKCP", "SYNCS", "SHFL", "BAR",
            "LDL", "STL", "ATOM", "RED", "CCTL", "UCGABAR", "MEMBAR", "ERRBAR", "LDSM", "BRA")


def main():
    tag, wanted = sys.argv[1], sys.argv[2:]
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "--dump-resource-usage", LIB], capture_output=True, text=True).stdout
    usage = {}
    for m in re.finditer(r"Function (\S+):\s*\n\s*(REG:\d+.*)", res):
        usage[m.group(1)] = m.group(2).strip()
    out = {}
    os.makedirs(os.path.join(ROOT, "profiles", "sass"), exist_ok=True)
    for blk in re.split(r"\n\s*Function : ", sass)[1:]:
        mangled = blk.split("\n", 1)[0].strip()
        short = next((w for w in wanted if re.search(r"\d+%s(E|I|P|v|N)" % re.escape(w), mangled) or mangled.endswith(w)), None)
        if wanted and not short:
            continue
        short = short or mangled
        ops = collections.Counter()
        n = 0
        for line in blk.splitlines():
            m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
            if not m:
                continue
            n += 1
            op = m.group(1)
            base = op.split(".")[0]
            if base in INTEREST:
                key = op if base in ("LDS", "STS", "LDG", "STG", "LDGSTS", "LDL", "STL", "UBLKCP", "BAR", "SHFL", "MUFU") else base
                ops[key] += 1
        out[short] = {"mangled": mangled, "instructions": n, "opcodes": dict(sorted(ops.items())),
                      "spill_instructions": sum(v for k, v in ops.items() if k.startswith(("LDL", "STL"))),
                      "resources": usage.get(mangled)}
        with open(os.path.join(ROOT, "profiles", "sass", short + ".sass"), "w") as fh:  # mnemonics only: the hex encodings double the size
            body = re.sub(r"[ \t]*/\* 0x[0-9a-f]{16} \*/", "", "Function : " + blk)
            fh.write("\n".join(l.rstrip() for l in body.splitlines() if l.strip()) + "\n")
    try:
        with open(os.path.join(ROOT, "viso_b200", ".libviso_ba.stamp")) as fh:
            digest = fh.read().strip()
    except Exception:
        digest = None
    path = os.path.join(ROOT, "profiles", "%s_sass_summary.json" % tag)
    with open(path, "w") as fh:
        json.dump({"source_digest": digest, "kernels": out}, fh, indent=1)
    for k, v in out.items():
        print(k, v["instructions"], "instr; spills", v["spill_instructions"], ";", v["resources"])
        print("   ", {a: b for a, b in v["opcodes"].items() if a.startswith(("D", "LDS", "LDGSTS", "UBLKCP", "MUFU"))})


if __name__ == "__main__":
    main()
