#!/usr/bin/env python
"""Distils ncu captures into profiles/<round>_ncu_constants.json — the profiler-only counters bench.py quotes
(DRAM traffic per launch, shared-memory pipe, tensor pipe ...), each with the sha of the kernel sources at
capture time so that bench.py can flag them stale when the kernel changes.

    python tools/ncu_constants.py k_score_fused=gpurun_out/r2a_fused.ncu-rep k_mlp_gemm=gpurun_out/r2a_ltv.ncu-rep \
        c3_build=gpurun_out/r2a_c3_launches.csv [--out profiles/r2_ncu_constants.json] [--csv-prefix profiles/r2a]

A `.ncu-rep` argument is read with `ncu -i ... --page raw --csv` (works without a GPU); a `.csv` argument is a
launch list written by `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv`.
"""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = "igaming-platform_b200/csrc/"
SOURCES = {
    "k_score_fused": [CSRC + "k_score.cu", CSRC + "velocity.cuh", CSRC + "igs_math.cuh"],
    "k_score_features": [CSRC + "k_score.cu", CSRC + "igs_math.cuh"],
    "k_mlp_gemm": [CSRC + "k_ltv.cu"],
    "k_partition": [CSRC + "k_partition.cu"],
    "c3_build": [CSRC + "k_velocity.cu", CSRC + "sortscan.cuh", CSRC + "onesweep.cuh", CSRC + "velocity.cuh"],
}
KEEP = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": "smem_wavefronts",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed": "smem_pipe_pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflicts",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "alu_pipe_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pipe_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "lts__t_sectors_srcunit_tex_op_read.sum": "l2_read_sectors_from_l1",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum": "global_load_sectors",
    "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum": "global_load_requests",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
}
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}


def sha16(paths):
    h = hashlib.sha256()
    for p in paths:
        with open(os.path.join(ROOT, p), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def num(v, unit):
    try:
        x = float(v.replace(",", ""))
    except ValueError:
        return None
    return x * UNIT.get(unit, 1.0)


def from_rep(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    launches = []
    for r in rows[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")]}
        for k, name in KEEP.items():
            if k in hdr:
                d[name] = num(r[hdr.index(k)], units[hdr.index(k)])
        launches.append(d)
    return launches, raw


def from_launch_csv(path):
    rows = list(csv.reader(open(path)))
    hdr = next(r for r in rows if r and r[0] == "ID")
    per = {}
    order = []
    for r in rows:
        if len(r) < len(hdr) or r[0] == "ID":
            continue
        d = dict(zip(hdr, r))
        key = (d["ID"], d["Kernel Name"])
        if key not in per:
            per[key] = {"kernel": d["Kernel Name"]}
            order.append(key)
        name = KEEP.get(d["Metric Name"])
        if name:
            per[key][name] = num(d["Metric Value"], d["Metric Unit"])
    return [per[k] for k in order]


def short(name):
    name = name.replace("<unnamed>::", "").replace("void ", "")
    return name.split("(")[0]


def main():
    out_path = os.path.join(ROOT, "profiles", "r2_ncu_constants.json")
    prefix = None
    items = []
    args = sys.argv[1:]
    while args:
        a = args.pop(0)
        if a == "--out":
            out_path = args.pop(0)
        elif a == "--csv-prefix":
            prefix = args.pop(0)
        else:
            items.append(a.split("=", 1))
    res = json.load(open(out_path)) if os.path.exists(out_path) else {}
    for key, path in items:
        src = SOURCES.get(key, [])
        rec = {"source": os.path.basename(path), "sources": src, "source_sha16": sha16(src)}
        if path.endswith(".ncu-rep"):
            launches, raw = from_rep(path)
            if prefix:
                with open(f"{prefix}_{key}_ncu_raw.csv", "w") as f:      # the selected columns only, per launch
                    w = csv.writer(f)
                    cols = ["kernel"] + list(KEEP.values())
                    w.writerow(cols)
                    for l in launches:
                        w.writerow([l.get(c, "") for c in cols])
                rec["source"] = os.path.basename(f"{prefix}_{key}_ncu_raw.csv")
            by = {}
            for l in launches:
                by.setdefault(short(l["kernel"]), []).append(l)
            rec["kernels"] = {}
            for k, ls in by.items():
                agg = {"launches": len(ls)}
                for name in KEEP.values():
                    vals = [l[name] for l in ls if l.get(name) is not None]
                    if vals:
                        agg[name] = sum(vals) / len(vals)
                rec["kernels"][k] = agg
            first = max(rec["kernels"].values(), key=lambda v: v.get("duration") or 0)    # the dominant kernel of the capture
            rec["dram_bytes"] = (first.get("dram_read") or 0) + (first.get("dram_write") or 0)
            rec["duration_ms"] = first.get("duration")
        else:
            launches = from_launch_csv(path)
            by = {}
            for l in launches:
                by.setdefault(short(l["kernel"]), []).append(l)
            rec["kernels"] = {}
            for k, ls in by.items():
                t = sum(l.get("duration") or 0 for l in ls)
                rd = sum(l.get("dram_read") or 0 for l in ls)
                wr = sum(l.get("dram_write") or 0 for l in ls)
                rec["kernels"][k] = {"launches": len(ls), "total_ms": t, "dram_read": rd, "dram_write": wr,
                                     "gbs_moved": (rd + wr) / (t * 1e-3) / 1e9 if t else None}
            rec["total_ms"] = sum(v["total_ms"] for v in rec["kernels"].values())
            if prefix:
                import shutil
                shutil.copy(path, f"{prefix}_{key}_launches.csv")
                rec["source"] = os.path.basename(f"{prefix}_{key}_launches.csv")
        res[key] = rec
    json.dump(res, open(out_path, "w"), indent=1)
    print("wrote", out_path, "keys:", sorted(res))


if __name__ == "__main__":
    main()
