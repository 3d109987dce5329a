#!/usr/bin/env python
"""Summarise `ncu --set full` reports (gpurun_out/*.ncu-rep) into small text files under profiles/.

    python scripts/ncu_summary.py gpurun_out/prof_x.ncu-rep [...] --out profiles/r1_ncu_full.md
    python scripts/ncu_summary.py --launches gpurun_out/launches_x.csv --out profiles/r1_launches.md

Runs here (no GPU): `ncu -i <rep> --page raw --csv` and `--page source --csv` read the report offline.
"""
from __future__ import annotations

import argparse
import csv
import io
import json
import subprocess
import sys
from collections import defaultdict
from pathlib import Path

RAW_KEYS = [
    "gpu__time_duration.sum",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "sm__cycles_active.avg", "sm__cycles_elapsed.max",
]


def ncu_csv(rep: Path, page: str):
    out = subprocess.run(["ncu", "-i", str(rep), "--page", page, "--csv"], capture_output=True, text=True).stdout
    lines = [l for l in out.splitlines() if l.startswith('"')]
    return list(csv.reader(io.StringIO("\n".join(lines))))


def to_bytes(val: str, unit: str) -> float:
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(val) * mult.get(unit, 1)


def summarise_full(rep: Path):
    rows = ncu_csv(rep, "raw")
    hdr, units = rows[0], rows[1]
    res = []
    for vals in rows[2:]:
        d = {h: (vals[i], units[i]) for i, h in enumerate(hdr)}
        res.append(d)
    return res


def stall_summary(rep: Path, top=12):
    """Aggregate the per-instruction sampling columns of the source page; list the hottest source lines."""
    rows = ncu_csv(rep, "source")
    while rows and rows[0][0] != "Address":
        rows = rows[1:]
    if not rows:
        return None
    hdr = rows[0]
    idx = {h: i for i, h in enumerate(hdr)}
    samp_col = "# Samples" if "# Samples" in idx else None
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = defaultdict(float)
    lines = []
    for r in rows[1:]:
        if len(r) != len(hdr):
            continue
        for c in stall_cols:
            try:
                tot[c] += float(r[idx[c]] or 0)
            except ValueError:
                pass
        if samp_col:
            try:
                s = float(r[idx[samp_col]] or 0)
            except ValueError:
                s = 0
            src = r[idx.get("Source", 1)] if "Source" in idx else ""
            lines.append((s, src.strip()[:110]))
    lines.sort(reverse=True)
    return {"stalls": sorted(tot.items(), key=lambda kv: -kv[1])[:10], "hot": lines[:top], "samp_col": samp_col}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("reports", nargs="*")
    ap.add_argument("--launches", default=None)
    ap.add_argument("--out", required=True)
    ap.add_argument("--title", default="")
    ap.add_argument("--traffic-json", default=None, help="update profiles/traffic.json: workload=report pairs")
    ap.add_argument("--workloads", nargs="*", default=[], help="workload name per report (for --traffic-json)")
    args = ap.parse_args()
    out = []
    if args.title:
        out.append(f"# {args.title}\n")
    if args.launches:
        rows = list(csv.DictReader(l for l in open(args.launches) if l.startswith('"')))
        out.append(f"Launch list: `{args.launches}` ({len(rows)} launches; `ncu --metrics gpu__time_duration.sum "
                   "--clock-control none`, cold-cache and serialised: compare shares, not absolutes)\n")
        agg = defaultdict(list)
        for r in rows:
            name = r["Kernel Name"].split("(")[0].replace("void <unnamed>::", "")
            agg[(name, r["Grid Size"], r["Block Size"])].append(float(r["Metric Value"]))
        total = sum(sum(v) for v in agg.values())
        out.append("| kernel | grid | block | launches | mean ns | min ns | share of listed time |")
        out.append("|---|---|---|---|---|---|---|")
        for (name, g, b), v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
            out.append(f"| `{name}` | {g} | {b} | {len(v)} | {sum(v) / len(v):.0f} | {min(v):.0f} | {100 * sum(v) / total:.1f} % |")
        out.append("")
    traffic = {}
    for k, rep in enumerate(args.reports):
        rep = Path(rep)
        for d in summarise_full(rep):
            out.append(f"## `{rep.name}`\n")
            out.append(f"kernel: `{d['Kernel Name'][0][:140]}`\n")
            out.append("| metric | value | unit |")
            out.append("|---|---|---|")
            for key in RAW_KEYS:
                if key in d and d[key][0] != "":
                    out.append(f"| {key} | {d[key][0]} | {d[key][1]} |")
            rd = to_bytes(*d["dram__bytes_read.sum"])
            wr = to_bytes(*d["dram__bytes_write.sum"])
            out.append(f"| **traffic = dram read + write** | {rd + wr:.0f} | byte |")
            out.append("")
            if k < len(args.workloads):
                # bench.py reports the number only for the per-GPU shape it was captured on
                sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
                from bench import WORKLOADS
                w = WORKLOADS[args.workloads[k]]
                traffic[args.workloads[k]] = {"per_gpu": {"B": w["B"], "Hq": w["Hq"], "Hkv": w["Hkv"]},
                                              "dram_bytes_per_launch": int(rd + wr), "capture": args.title or rep.name}
        st = stall_summary(rep)
        if st and st["stalls"]:
            out.append("warp-state samples by stall reason (source page, summed over instructions):\n")
            for c, v in st["stalls"]:
                if v > 0:
                    out.append(f"- {c}: {v:.0f}")
            out.append("")
        if st and st["hot"]:
            out.append(f"hottest source lines by `{st['samp_col']}`:\n")
            out.append("```")
            for s, src in st["hot"]:
                out.append(f"{s:8.0f}  {src}")
            out.append("```\n")
    Path(args.out).write_text("\n".join(out) + "\n")
    if args.traffic_json and traffic:
        p = Path(args.traffic_json)
        cur = json.loads(p.read_text()) if p.exists() else {}
        cur.update(traffic)
        p.write_text(json.dumps(cur, indent=1, sort_keys=True) + "\n")
    print(f"wrote {args.out}", file=sys.stderr)


if __name__ == "__main__":
    main()
