#!/usr/bin/env python
"""Turns the ncu outputs a GPU visit leaves in gpurun_out/ into the tracked summaries under profiles/.

    python scripts/summarize_profile.py <tag>      # e.g. r1a  -> profiles/<tag>_launches.md, <tag>_kernels.csv

Inputs (all optional): gpurun_out/launches.csv (the `--metrics gpu__time_duration.sum` pass),
gpurun_out/prof_*.ncu-rep (`--set full` captures), gpurun_out/bench.json.
"""
from __future__ import annotations

import collections
import csv
import glob
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GO = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

RAW_METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active",
    "smsp__warps_eligible.avg.per_cycle_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
]
STALL_RE = re.compile(r"^smsp__average_warps?_(?:latency_)?issue_stalled_(.+?)(?:_per_issue_active)?\.(?:ratio|pct)$")


def short_name(full: str) -> str:
    s = re.sub(r"^void ", "", full)
    s = s.replace("q3a::", "").replace("__nv_bfloat16", "bf16").replace("__half", "f16")
    s = re.sub(r"\((?:const |CUtensorMap_st|float|int|unsigned|long|T\d).*$", "", s)
    return s


def launches(tag: str) -> None:
    """gpurun_out/launches.csv: one ncu pass over the launches of ONE bench step with
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --print-units base."""
    path = os.path.join(GO, "launches.csv")
    whole_run = os.path.join(GO, "launches_all.csv")       # every launch of the run: one device-resident step is cut out below
    if os.path.exists(whole_run):
        path = whole_run
    if not os.path.exists(path):
        return
    rows = list(csv.reader(open(path, newline="")))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr = rows[hi]
    kn, mn, mv = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    gs, bs, idc = hdr.index("Grid Size"), hdr.index("Block Size"), hdr.index("ID")
    per_launch: dict = collections.OrderedDict()
    for r in rows[hi + 1:]:
        if len(r) <= mv or not r[idc].isdigit():
            continue
        d = per_launch.setdefault(r[idc], {"key": (short_name(r[kn]), r[gs], r[bs])})
        try:
            d[r[mn]] = float(r[mv].replace(",", ""))
        except ValueError:
            pass
    if path == whole_run:
        # a device-resident step starts with its single mel_frames_kernel launch: keep [2nd such launch, 3rd) -- past the
        # first (cold) step, before the host-pointer steps (which launch the mel kernel once per upload part)
        ids = list(per_launch)
        starts = [i for i, k in enumerate(ids) if per_launch[k]["key"][0].startswith("mel_frames_kernel")]
        if len(starts) >= 3:
            keep = ids[starts[1]:starts[2]]
            per_launch = collections.OrderedDict((k, per_launch[k]) for k in keep)
    agg: dict = collections.OrderedDict()
    total = 0.0
    for d in per_launch.values():
        a = agg.setdefault(d["key"], [0, 0.0, 0.0])
        a[0] += 1
        a[1] += d.get("gpu__time_duration.sum", 0.0)
        a[2] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        total += d.get("gpu__time_duration.sum", 0.0)
    n = len(per_launch)
    with open(os.path.join(PROF, f"{tag}_launches.md"), "w") as f:
        f.write(f"# ncu launch list — {tag}\n\n")
        f.write("`ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none` over one "
                "step of `bench.py` (64 x 30-s clips, ASR-0.6B, bf16 operands).  Times are cold-cache and serialised: "
                "compare shares.  DRAM = read + write bytes of all launches of the row.\n\n")
        f.write(f"{n} launches, {total / 1e6:.3f} ms in total.\n\n")
        f.write("| kernel | grid | block | launches | total ms | share | mean us | DRAM MB | GB/s |\n|---|---|---|---:|---:|---:|---:|---:|---:|\n")
        for (name, g, b), (cnt, ns, by) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{name}` | {g} | {b} | {cnt} | {ns / 1e6:.3f} | {100 * ns / total:.1f}% | {ns / cnt / 1e3:.1f} | "
                    f"{by / 1e6:.1f} | {by / max(ns, 1.0):.0f} |\n")
    traffic = {}
    for (name, g, b), (cnt, ns, by) in agg.items():
        fam = name.split("<")[0]
        t = traffic.setdefault(fam, {"launches": 0, "dram_bytes": 0.0, "ns": 0.0})
        t["launches"] += cnt; t["dram_bytes"] += by; t["ns"] += ns
    with open(os.path.join(PROF, f"{tag}_traffic.json"), "w") as f:
        json.dump({"source": f"profiles/{tag}_launches.md (ncu dram__bytes_read.sum + dram__bytes_write.sum, one bench step)",
                   "kernels": traffic}, f, indent=1)
    g = traffic.get("gemm_tc_kernel")
    if g and g["launches"] and g["dram_bytes"] > 0:
        with open(os.path.join(PROF, "roofline_traffic.json"), "w") as f:       # read by bench.py -> roofline.traffic
            out = {"gemm_tc_kernel_dram_bytes_per_launch": g["dram_bytes"] / g["launches"], "launches_per_step": g["launches"],
                   "source": f"profiles/{tag}_traffic.json"}
            # hash of csrc/ as it was on the GPU box when the launch list was taken (the gpurun script writes it next to
            # the csv); bench.py reports traffic only while the sources it runs still hash to this
            hp = os.path.join(GO, "csrc_sha256.txt")
            out["csrc_sha256"] = open(hp).read().strip() if os.path.exists(hp) else None
            for fam in ("attention_kernel", "mel_frames_kernel"):
                t = traffic.get(fam)
                if t and t["launches"]:
                    out[f"{fam}_dram_bytes_per_launch"] = t["dram_bytes"] / t["launches"]
            json.dump(out, f, indent=1)
    print("wrote", f"profiles/{tag}_launches.md", f"profiles/{tag}_traffic.json")


def full_reports(tag: str) -> None:
    out_rows = []
    for rep in sorted(glob.glob(os.path.join(GO, "prof_*.ncu-rep"))):
        r = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--print-units", "base"], capture_output=True, text=True)
        if r.returncode != 0:
            print("ncu -i failed on", rep, r.stderr[:200])
            continue
        rows = list(csv.reader(io.StringIO(r.stdout)))
        hdr = rows[0]
        col = {h: i for i, h in enumerate(hdr)}
        for row in rows[2:]:
            d = {"report": os.path.basename(rep), "kernel": short_name(row[col["Kernel Name"]])}
            for m in RAW_METRICS:
                if m in col:
                    d[m] = row[col[m]]
            # the three largest warp-stall reasons (cycles a warp waits per instruction it issues)
            stalls = []
            for h, i in col.items():
                mm = STALL_RE.match(h)
                if mm and "not_issued" not in h:
                    try:
                        stalls.append((float(row[i].replace(",", "")), mm.group(1)))
                    except ValueError:
                        pass
            d["top_stalls"] = "; ".join(f"{n} {v:.2f}" for v, n in sorted(stalls, reverse=True)[:3])
            out_rows.append(d)
    if not out_rows:
        return
    keys = ["report", "kernel"] + RAW_METRICS + ["top_stalls"]
    with open(os.path.join(PROF, f"{tag}_kernels.csv"), "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=keys)
        w.writeheader()
        for d in out_rows:
            w.writerow({k: d.get(k, "") for k in keys})
    print("wrote", f"profiles/{tag}_kernels.csv")


def bench(tag: str) -> None:
    for name in ("bench.json", "bench_ref.json", "r2b_bench.json", "r2b_bench_ref.json"):
        p = os.path.join(GO, name)
        if os.path.exists(p) and os.path.getsize(p) > 0:
            with open(p) as f:
                txt = f.read().strip().splitlines()[-1]
            try:
                json.loads(txt)
            except Exception:
                continue
            with open(os.path.join(PROF, f"{tag}_{name}"), "w") as f:
                f.write(txt + "\n")
            print("wrote", f"profiles/{tag}_{name}")


if __name__ == "__main__":
    tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
    os.makedirs(PROF, exist_ok=True)
    launches(tag)
    full_reports(tag)
    bench(tag)
