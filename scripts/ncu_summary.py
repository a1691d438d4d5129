"""Summarise .ncu-rep captures (ncu --set full) into a small JSON for profiles/: per kernel the duration, launch shape,
DRAM traffic, pipe utilisations and stall reasons.  Usage: python scripts/ncu_summary.py out.json rep1.ncu-rep [rep2 ...]"""
import csv
import io
import json
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active",
        "sm__ops_path_tensor_src_fp64.avg.pct_of_peak_sustained_elapsed",
        "sm__ops_path_tensor_src_int8.avg.pct_of_peak_sustained_elapsed",
        "sm__ops_path_tensor_op_utcimma_src_int8_sparsity_off.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def summarise(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    out = {}
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        d = {}
        for h, u, v in zip(hdr, units, r):
            if h in KEEP or (h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio")
                             and v not in ("", "n/a") and float(v) >= 0.2):
                d[h] = f"{v} {u}".strip()
        key = f"{rep.split('/')[-1].replace('.ncu-rep', '')}:{name[:90]}"
        k, i = key, 1
        while k in out:
            i += 1
            k = f"{key}#{i}"
        out[k] = d
    return out


if __name__ == "__main__":
    res = {}
    for rep in sys.argv[2:]:
        res.update(summarise(rep))
    json.dump(res, open(sys.argv[1], "w"), indent=1)
    print(json.dumps(res, indent=1)[:3000])
