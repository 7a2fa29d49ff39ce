#!/usr/bin/env python
"""Register / spill / shared-memory report of every kernel (`nvcc -Xptxas -v`, cross-compiled for sm_100a; no GPU needed).
Writes profiles/ptxas_report.md: one row per kernel entry, spills flagged.

    python tools/ptxas_report.py
"""
from __future__ import annotations

import glob
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def demangle(names):
    try:
        out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True, check=True).stdout.splitlines()
        return out if len(out) == len(names) else names
    except Exception:  # noqa: BLE001
        return names


def main() -> int:
    rows = []
    for src in sorted(glob.glob(os.path.join(ROOT, "csrc", "kernels", "*.cu"))):
        cmd = ["nvcc", "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-Xptxas", "-v",
               "-I", os.path.join(ROOT, "include"), "-c", src, "-o", os.devnull]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            print(r.stderr[-2000:], file=sys.stderr)
            return 1
        cur = None
        for line in r.stderr.splitlines():
            m = re.search(r"Compiling entry function '([^']+)' for 'sm_100a'", line)
            if m:
                cur = {"file": os.path.basename(src), "name": m.group(1), "regs": 0, "spill_st": 0, "spill_ld": 0, "stack": 0,
                       "smem": 0}
                rows.append(cur)
                continue
            if cur is None:
                continue
            m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m:
                cur["stack"], cur["spill_st"], cur["spill_ld"] = map(int, m.groups())
            m = re.search(r"Used (\d+) registers", line)
            if m:
                cur["regs"] = int(m.group(1))
                sm = re.search(r"(\d+) bytes smem", line)
                cur["smem"] = int(sm.group(1)) if sm else 0
    names = demangle([r["name"] for r in rows])
    out = ["# ptxas report (`nvcc -Xptxas -v`, sm_100a; `tools/ptxas_report.py`)", "",
           "Static shared memory only (dynamic shared memory is requested at launch). A kernel with spill bytes is flagged.", "",
           "| file | kernel | registers | stack B | spill st / ld B | static smem B |", "|---|---|---|---|---|---|"]
    spills = 0
    for r, n in zip(rows, names):
        n = re.sub(r"\(.*$", "", n)
        flag = " **spills**" if (r["spill_st"] or r["spill_ld"]) else ""
        spills += bool(flag)
        out.append(f"| {r['file']} | `{n}` | {r['regs']} | {r['stack']} | {r['spill_st']} / {r['spill_ld']}{flag} | {r['smem']} |")
    out += ["", f"{len(rows)} kernel entries, {spills} with register spills."]
    with open(os.path.join(ROOT, "profiles", "ptxas_report.md"), "w") as f:
        f.write("\n".join(out) + "\n")
    print(f"{len(rows)} kernel entries, {spills} with spills -> profiles/ptxas_report.md")
    return 0


if __name__ == "__main__":
    sys.exit(main())
