"""bench.py --workload sdf: compute_sdf - squared distance, sign, normal and closest point of P points to a
triangle mesh of F faces (TorchSDF's kernel, thirdparty/TorchSDF/torchsdf/csrc/unbatched_triangle_distance_cuda.cu:176-254).
One step = one compute_sdf call over all points of this rank (points sharded by contiguous ranges, mesh replicated, no
collective).  When oracle/_ref/torchsdf_ref.so exists (the reference's kernel compiled unmodified for sm_100a) it is timed on
the same inputs in the same run: that kernel recompiled is the bar this row has to beat."""
import ctypes
import os
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
METRIC = "mesh SDF point queries/sec"
FLOP_PER_PAIR = 61   # 9 sub, 3 edge projections (5 + 1 div each), 3 edge-normal tests (5 each), plane projection 11, |p-c|^2 8


def _inputs(P_total, subdiv, rank, world):
    from compliancedex_b200.distributed import shard_range
    from compliancedex_b200.synthetic import icosphere_faces
    fv = icosphere_faces(subdiv, 0.1, np.float32)                       # object-sized sphere (radius 0.1 m), 20 * 4^subdiv faces
    lo, hi = shard_range(P_total, rank, world)
    g = torch.Generator().manual_seed(1234)
    p = ((torch.rand(P_total, 3, generator=g) * 2 - 1) * 0.2)[lo:hi].contiguous()
    return p, torch.from_numpy(fv)


def run_bench_sdf(args, rank, world, local, ClockSampler, barrier, max_over_ranks, fp64_pipe_peaks, ncu_traffic=lambda *a: None):
    from compliancedex_b200 import _lib
    from compliancedex_b200.sdf import compute_sdf
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    P_total, subdiv = args.points, args.subdiv
    ph, fh = _inputs(P_total, subdiv, rank, world)
    p, fv = ph.to(dev), fh.to(dev)
    P, F = p.shape[0], fv.shape[0]
    lib = _lib.load()
    ctx = _lib.context(local)

    from compliancedex_b200.sdf import set_search_mode
    set_search_mode(args.sdf_mode)

    def step():
        return compute_sdf(p, fv)

    for _ in range(args.warmup):
        out = step()
    sampler = ClockSampler(local)
    l0 = lib.cdx_launch_count(ctx)
    barrier(world)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        out = step()
    e1.record()
    barrier(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    launches = lib.cdx_launch_count(ctx) - l0

    # e2e: host points in, all four outputs back on the host
    pp = ph.pin_memory()
    hout = [torch.empty(P).pin_memory(), torch.empty(P, dtype=torch.int32).pin_memory(), torch.empty(P, 3).pin_memory(),
            torch.empty(P, 3).pin_memory()]

    def step_e2e():
        o = compute_sdf(pp.to(dev, non_blocking=True), fv)
        for h, d in zip(hout, o):
            h.copy_(d, non_blocking=True)
        torch.cuda.synchronize()

    step_e2e()
    barrier(world)
    e0.record()
    for _ in range(args.steps):
        step_e2e()
    e1.record()
    barrier(world)
    ms_e2e = max_over_ranks(e0.elapsed_time(e1), world) / args.steps

    # the exhaustive scan (the reference's algorithm on this repo's kernels) for comparison when the culled search was timed
    exhaustive_ms = None
    if args.sdf_mode != 1:
        set_search_mode(1)
        out1 = step()
        e0.record()
        for _ in range(args.steps):
            out1 = step()
        e1.record()
        torch.cuda.synchronize()
        exhaustive_ms = e0.elapsed_time(e1) / args.steps
        assert all(torch.equal(a, b) for a, b in zip(out, out1)), "culled search and exhaustive scan disagree"
        set_search_mode(args.sdf_mode)
    # the reference's own kernel, recompiled for sm_100a, same inputs, same run
    ref = None
    try:
        from oracle import ref_torchsdf
        if ref_torchsdf.load() is not None:
            for _ in range(2):
                r = ref_torchsdf.compute_sdf(p, fv)
            e0.record()
            for _ in range(args.steps):
                r = ref_torchsdf.compute_sdf(p, fv)
            e1.record()
            torch.cuda.synchronize()
            ref_ms = e0.elapsed_time(e1) / args.steps
            ref = {"ms_per_step": ref_ms, "queries_per_s": P / (ref_ms * 1e-3), "speedup_of_cdx": ref_ms / ms,
                   "what": "unbatched_triangle_distance_forward_cuda from the reference's unmodified sources, nvcc sm_100a "
                           "(oracle/_ref/torchsdf_ref.so), incl. the torch.zeros of its Python wrapper",
                   "dist_bit_exact": float((r[0] == out[0]).float().mean()), "closest_bit_exact": float((r[3] == out[3]).all(dim=1).float().mean()),
                   "sign_equal": float((r[1] == out[1]).float().mean())}
    except Exception as e:  # the checker is optional on the GPU box
        ref = {"unavailable": str(e)[:200]}
    if rank != 0:
        return None
    pipes = fp64_pipe_peaks()
    pairs = float(P) * F
    achieved = FLOP_PER_PAIR * pairs / (ms * 1e-3) / 1e12
    peak = pipes.get("ffma_tflops")
    return {
        "metric": METRIC, "value": P * world / (ms * 1e-3), "unit": "queries/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "sdf: compute_sdf of %d uniform points in [-0.2,0.2]^3 against an icosphere of %d faces (radius "
                               "0.1), float32, sharded by contiguous ranges over %d GPU(s)" % (P_total, F, world),
                   "pairs_per_step": pairs, "l2_policy": "compute-bound: 12 B in / 44 B out per point against %d face "
                   "evaluations; the %d KB of face records stay L2-resident by design" % (F, F * 160 // 1024)},
        "e2e": {"value": P * world / (ms_e2e * 1e-3), "unit": "queries/s", "h2d_bytes_per_step": int(P * 12),
                "d2h_bytes_per_step": int(P * (4 + 4 + 12 + 12))},
        "gpu_launches": int(launches), "clocks": clocks,
        "roofline": {"bound": "fp32", "kernel": "sdf_pair_kernel<float> (FP32 FMA pipe; bulk-TMA staged face records)",
                     "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
                     "peak_source": "FFMA rate measured in this run by csrc/cdx_pipebench (MEASURED_PEAKS.json has no FP32 figure)",
                     "flop_per_pair": FLOP_PER_PAIR, "pairs_per_launch": pairs, "pairs_per_s": pairs / (ms * 1e-3),
                     "traffic": ncu_traffic("r01_ncu_sdf_summary.json", "r01_prof_sdf_v2") if (P_total, subdiv, world) == (1_000_000, 5, 1) else None},
        "reference_kernel": ref,
        "search": {"mode": {0: "auto (culled for F >= 512 and P >= 128 x SMs)", 1: "exhaustive", 2: "culled"}[args.sdf_mode],
                   "exhaustive_scan_ms_per_step": exhaustive_ms,
                   "note": "roofline.achieved counts ALL P x F pairs (the work the reference does); the culled search visits only "
                           "the face tiles whose bounding sphere can still hold the closest face, and returns the same face bit for bit"},
    }


def cpu_sdf(n_points, subdiv, threads):
    """numpy restatement of the reference kernel (oracle/sdf_oracle.py, kind "port": the reference has no CPU path for this
    op) on a sample of the workload -> queries/s, seconds"""
    from oracle import sdf_oracle as so
    fv = so.icosphere(subdiv, 0.1, np.float32)
    rng = np.random.default_rng(0)
    p = ((rng.random((n_points, 3)) * 2 - 1) * 0.2).astype(np.float32)
    t0 = time.perf_counter()
    so.compute_sdf(p, fv)
    dt = time.perf_counter() - t0
    return n_points / dt, dt
