#!/usr/bin/env python
"""Per-CUDA-source-line view of an `ncu --set full --import-source on` report (kernels built with -lineinfo):

    python scripts/ncu_lines.py gpurun_out/prof_x.ncu-rep [--top 40]

Aggregates the warp-state samples and executed instructions of the SASS rows by the CUDA line they map to.
Runs offline (no GPU): reads `ncu -i <rep> --page source --csv --print-source sass,cuda`.
"""
from __future__ import annotations

import argparse
import csv
import io
import subprocess
from collections import defaultdict
from pathlib import Path


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep", type=Path)
    ap.add_argument("--top", type=int, default=40)
    a = ap.parse_args()
    txt = subprocess.run(["ncu", "-i", str(a.rep), "--page", "source", "--csv", "--print-source", "sass,cuda"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    samples = defaultdict(int)
    insts = defaultdict(int)
    text = {}
    cur_file, hdr = "?", None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = Path(r[1]).name
            hdr = None
            continue
        if r[0] == "Line No":
            hdr = r
            i_s = hdr.index("# Samples")
            i_i = hdr.index("Instructions Executed")
            continue
        if hdr is None or len(r) <= i_s:
            continue
        if not r[0].strip():      # SASS rows carry no line number; the CUDA rows already aggregate them
            continue
        key = (cur_file, r[0])
        text.setdefault(key, r[1].strip())
        try:
            samples[key] += int(r[i_s] or 0)
            insts[key] += int(r[i_i] or 0)
        except ValueError:
            pass
    tot_s, tot_i = sum(samples.values()) or 1, sum(insts.values()) or 1
    print(f"total samples {tot_s}, warp instructions {tot_i}")
    print(f"{'samples':>8} {'%':>5} {'inst %':>6}  file:line  source")
    for key in sorted(samples, key=lambda k: -samples[k])[: a.top]:
        print(f"{samples[key]:8d} {100 * samples[key] / tot_s:5.1f} {100 * insts[key] / tot_i:6.1f}  {key[0]}:{key[1]}  {text[key][:110]}")


if __name__ == "__main__":
    main()
