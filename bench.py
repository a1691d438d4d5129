#!/usr/bin/env python
"""bench.py - hot-path throughput of compliancedex_b200 on B200 (contract in the task brief).

BASELINE.json's metric has two halves, "GPIS mean+var+grad queries/sec and pregrasp candidate-iters/sec at 1/2/4/8 B200".
The default run (`--workload auto`) prints ONE JSON line whose top-level keys are the first half on the config that spells
out the scaling - gpis_c3 (configs[2]: synthetic GPIS N=4096, 10 M queries, STRONG scaling over the ranks) - and whose
`secondary` object carries, each with its own value / e2e / roofline / cpu_baseline / clocks:
  pregrasp_c5  configs[4]: Allegro on hammer, 65,536 candidates x 500 iterations in total split over the ranks, NCCL arg-min
  pregrasp_c2  configs[1]: Leap on mug, 1,024 candidates x 200 iterations          (N = 1 only)
  fit_c4       configs[3]: GPIS fit (blocked Cholesky) at N = 16,384                (N = 1 only)
  gpis_c1      configs[0]: banana state, 64^3 query grid                            (N = 1 only)
`--workload <name>` runs one of them (or `sdf`) as the primary line.

`--impl reference` times the reference's own CPU implementation of the same workloads on the host cores (CUDA hidden from the
process): its unmodified Python from oracle/_ref/reference_py (kind "reference"; oracle/build_ref_py.py) or, if that copy is
missing, the oracle port (kind "port"), each step a bounded sample.  Under torchrun only rank 0 works.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = ["gpis_c3", "pregrasp_c5", "pregrasp_c2", "fit_c4", "gpis_c1", "sdf"]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cdx", choices=["cdx", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto"] + WORKLOADS)
    ap.add_argument("--secondary", default="auto", help="auto | none | comma-separated workloads reported under `secondary`")
    ap.add_argument("--queries", type=int, default=10_000_000)
    ap.add_argument("--ntrain", type=int, default=4096)
    ap.add_argument("--candidates", type=int, default=1024, help="pregrasp_c2: candidates per GPU")
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--hand", default="leap", choices=["leap", "allegro"])
    ap.add_argument("--object", default="mug", help="pregrasp_c2: shipped GPIS state (mug, hammer, lego, ...)")
    ap.add_argument("--c5-candidates", type=int, default=65536, help="pregrasp_c5: candidates in TOTAL over all ranks")
    ap.add_argument("--c5-iters", type=int, default=500)
    ap.add_argument("--c5-object", default="hammer", choices=["hammer", "lego", "coffeebottle"])
    ap.add_argument("--fit-n", type=int, default=16384)
    ap.add_argument("--points", type=int, default=1_000_000, help="sdf workload: query points")
    ap.add_argument("--subdiv", type=int, default=5, help="sdf workload: icosphere subdivisions (20 * 4^n faces)")
    ap.add_argument("--sdf-mode", type=int, default=0, help="sdf workload: 0 auto (culled search), 1 exhaustive scan, 2 culled")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--var-mode", type=int, default=0,
                    help="gpis_c3: 0 auto (tcgen05 INT8 slices), -1 FP64 tensor MMA, 4..6 INT8 with that many slices")
    ap.add_argument("--ncu-summary-c3", default="r02_ncu_gpis_c3_summary.json")
    ap.add_argument("--ncu-summary-pg", default="r02_ncu_pregrasp_summary.json")
    return ap.parse_args()


def secondary_list(args, world):
    if args.workload != "auto" and args.secondary == "auto":
        return []
    if args.secondary == "none":
        return []
    if args.secondary != "auto":
        return [w for w in args.secondary.split(",") if w in WORKLOADS]
    return ["pregrasp_c5"] + (["pregrasp_c2", "fit_c4", "gpis_c1"] if world == 1 else [])


def run_cdx(workload, args, rank, world, local):
    """-> JSON line dict on rank 0 (None elsewhere), cpu_baseline included"""
    import bench_common as bc
    if workload in ("fit_c4", "gpis_c1") and world > 1:
        raise SystemExit("%s is a single-GPU workload" % workload)
    if workload == "gpis_c3":
        from bench_gpis import run_gpis_c3
        line = run_gpis_c3(args, rank, world, local)
    elif workload == "gpis_c1":
        from bench_gpis import run_gpis_c1
        line = run_gpis_c1(args, rank, world, local)
    elif workload == "fit_c4":
        from bench_gpis import run_fit_c4
        line = run_fit_c4(args, rank, world, local)
    elif workload in ("pregrasp_c2", "pregrasp_c5"):
        from bench_pregrasp import run_pregrasp
        line = run_pregrasp(args, rank, world, local, workload)
    else:
        from bench_sdf import cpu_sdf, run_bench_sdf
        line = run_bench_sdf(args, rank, world, local, bc.ClockSampler, bc.barrier, bc.max_over_ranks, bc.pipe_peaks, bc.ncu_traffic)
        if rank == 0 and not args.no_cpu_baseline:
            v, dt = cpu_sdf(1000, args.subdiv, 1)
            line["cpu_baseline"] = {"value": v, "unit": "queries/s", "cores": 1, "kind": "port",
                                    "sample": "1000 of the points against the same mesh through the numpy port of the reference "
                                              "kernel (the reference has no CPU path for this op), %.1f s" % dt}
        return line
    if rank == 0 and not args.no_cpu_baseline:
        extra = ["--ntrain", str(args.ntrain), "--queries", str(args.queries), "--candidates", str(args.candidates), "--iters",
                 str(args.iters), "--hand", args.hand, "--object", args.object, "--c5-candidates", str(args.c5_candidates),
                 "--c5-iters", str(args.c5_iters), "--c5-object", args.c5_object, "--fit-n", str(args.fit_n)]
        line["cpu_baseline"] = bc.cpu_baseline_subprocess(workload, extra)
    return line


def run_reference(workload, args, threads, world):
    import numpy as np
    if workload == "gpis_c3":
        from bench_gpis import ref_gpis_c3
        return ref_gpis_c3(args, threads, world)
    if workload == "gpis_c1":
        from bench_gpis import ref_gpis_c1
        return ref_gpis_c1(args, threads, world)
    if workload == "fit_c4":
        from bench_gpis import ref_fit_c4
        return ref_fit_c4(args, threads, world)
    if workload in ("pregrasp_c2", "pregrasp_c5"):
        from bench_pregrasp import ref_pregrasp
        return ref_pregrasp(args, threads, world, workload)
    from bench_sdf import METRIC, cpu_sdf
    vals = [cpu_sdf(1000, args.subdiv, 1) for _ in range(args.warmup + args.steps)][args.warmup:]
    v, ms = float(np.mean([a for a, _ in vals])), float(np.mean([b for _, b in vals])) * 1e3
    return {"impl": "reference", "metric": METRIC, "value": v, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "sdf: 1000-point sample against the icosphere of %d faces; the reference has no CPU "
                                   "implementation of this op, so this is the numpy port" % (20 * 4 ** args.subdiv)},
            "cpu_baseline": {"value": v, "unit": "queries/s", "cores": 1, "kind": "port", "sample": "1000 points per step"},
            "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main():
    args = parse()
    if args.impl == "reference":
        # the reference arm is a CPU measurement: hide the GPUs before torch is imported (oracle/ref_harness.py then forces
        # the reference's hard-coded .cuda() calls onto the host), no process group, rank 0 only
        os.environ["CUDA_VISIBLE_DEVICES"] = ""
        rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
        if rank != 0:
            return
        threads = os.cpu_count() or 1
        primary = "gpis_c3" if args.workload == "auto" else args.workload
        line = run_reference(primary, args, threads, world)
        sec = {}
        for w in secondary_list(args, world):
            one = argparse.Namespace(**vars(args))
            one.steps, one.warmup = 1, 0
            try:
                sec[w] = run_reference(w, one, threads, world)
            except Exception as e:  # noqa: BLE001
                sec[w] = {"error": repr(e)}
        if sec:
            line["secondary"] = sec
        print(json.dumps(line), flush=True)
        return

    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the cdx path has no CPU fallback")
    import bench_common as bc
    rank, world, local = bc.dist_setup()
    primary = "gpis_c3" if args.workload == "auto" else args.workload
    line = run_cdx(primary, args, rank, world, local)
    sec = {}
    for w in secondary_list(args, world):
        try:
            sec[w] = run_cdx(w, args, rank, world, local)
        except Exception as e:  # noqa: BLE001 - a failing secondary block must not lose the primary line
            if world > 1:
                raise
            sec[w] = {"error": repr(e)}
    if rank == 0:
        if sec:
            line["secondary"] = sec
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
