#!/usr/bin/env python
"""Turns one `ncu --set full` report into the small JSON that bench.py loads as `executed.from_capture`
(profiles/r<round>_primary_capture.json): the counters of ONE launch of the dominant kernel, with the commit and the
workload they were taken at — so the bench record can show them without passing them off as measured in that run.

    python tools/ncu_capture_json.py gpurun_out/prof.ncu-rep profiles/r2_primary_capture.json \
        --commit $(git rev-parse --short HEAD) --workload "1024 spheres 3840x2160 depth 1" [--launch 0]
"""
import argparse
import csv
import io
import json
import subprocess

WANT = {
    "gpu__time_duration.sum": "kernel_us",
    "smsp__inst_executed.sum": "warp_instructions_per_launch",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_slots_busy_pct",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "active_lanes_per_instruction",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "pipe_fma_pct",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "pipe_alu_pct",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active": "pipe_xu_pct",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active": "pipe_lsu_pct",
    "dram__bytes_read.sum": "dram_bytes_read",
    "dram__bytes_write.sum": "dram_bytes_written",
    "launch__registers_per_thread": "registers_per_thread",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum": "thread_fadd",
    "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum": "thread_fmul",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum": "thread_ffma",
}
UNIT = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("out")
    ap.add_argument("--commit", required=True)
    ap.add_argument("--workload", required=True)
    ap.add_argument("--launch", type=int, default=0, help="which profiled launch of the report to take")
    args = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", args.report, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    row = body[args.launch]
    out = {"source": "ncu --set full --clock-control none, one launch; NOT measured in the bench run that prints this",
           "commit": args.commit, "workload": args.workload, "kernel": row[hdr.index("Kernel Name")]}
    for metric, key in WANT.items():
        if metric in hdr:
            i = hdr.index(metric)
            try:
                out[key] = float(row[i].replace(",", "")) * UNIT.get(units[i], 1.0)
            except ValueError:
                pass
    if "dram_bytes_read" in out and "dram_bytes_written" in out:
        out["dram_bytes_per_launch"] = out["dram_bytes_read"] + out["dram_bytes_written"]
    if all(k in out for k in ("thread_fadd", "thread_fmul", "thread_ffma")):
        out["executed_fp32_flop_per_launch"] = out["thread_fadd"] + out["thread_fmul"] + 2.0 * out["thread_ffma"]
    with open(args.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
