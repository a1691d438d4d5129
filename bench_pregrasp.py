"""bench.py's pregrasp workloads: ProbabilisticGraspOptimizer.optimize through the drop-in class.
  pregrasp_c2 (BASELINE.json configs[1]): Leap hand on the mug GPIS, 1,024 candidates x 200 iterations per GPU (weak scaling).
  pregrasp_c5 (configs[4]): Allegro hand on the hammer (or lego / coffeebottle) GPIS, 65,536 candidates x 500 iterations in
                            TOTAL, split over the ranks by contiguous ranges (strong scaling), NCCL arg-min at the end.
One "step" = one whole optimisation (init + all iterations + arg-min).  The only cross-rank traffic is the final 16-byte
all-gather for the arg-min; candidate b of the global batch is the same candidate for every rank count."""
import ctypes
import os
import time

import numpy as np
import torch

from bench_common import measured_peaks, ncu_traffic, pipe_peaks, reference_kind, timed

ROOT = os.path.dirname(os.path.abspath(__file__))
METRIC = "pregrasp candidate-iters/sec"


def workload_name(tag, hand, obj, N, total, iters, world, weak):
    split = "%d candidates per GPU" % total if weak else "%d candidates in total split over the GPUs" % total
    return ("%s: ProbabilisticGraspOptimizer (optimize_target, optimize_palm), %s hand on the %s GPIS state (N=%d), %s x %d "
            "iterations" % (tag, hand.capitalize(), obj, N, split, iters))


def _candidates(hand, obj, total):
    from compliancedex_b200 import load_hand, state_path
    from compliancedex_b200.synthetic import surface_center, synthetic_candidates
    doc = load_hand(hand)
    d = np.load(state_path(f"{obj}_state"))
    center = surface_center(d["X1"], d["y1"])
    q0, T0, k0, palm = synthetic_candidates(doc, total, center, seed=100)
    return doc, q0, T0, k0, palm, int(d["X1"].shape[0])


def run_pregrasp(args, rank, world, local, tag):
    """tag 'pregrasp_c2': every rank optimises `--candidates` candidates of its own (global index rank * B + local);
    tag 'pregrasp_c5': `--c5-candidates` candidates in total, rank r takes the r-th contiguous range."""
    from compliancedex_b200 import GPIS, ProbabilisticGraspOptimizer, _lib, state_path
    from compliancedex_b200.distributed import global_argmin, shard_range
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    weak = tag == "pregrasp_c2"
    if weak:
        hand, obj, iters, total = args.hand, args.object, args.iters, args.candidates * world
        lo, hi = rank * args.candidates, (rank + 1) * args.candidates
    else:
        hand, obj, iters, total = "allegro", args.c5_object, args.c5_iters, args.c5_candidates
        lo, hi = shard_range(total, rank, world)
    doc, q0, T0, k0, palm, N = _candidates(hand, obj, total)
    q0, T0, k0, palm = q0[lo:hi].contiguous(), T0[lo:hi].contiguous(), k0[lo:hi].contiguous(), palm[lo:hi].copy()
    B = hi - lo
    gp = GPIS(0.08, 1.0)
    gp.load_state_data(state_path(f"{obj}_state"), device=dev)
    opt = ProbabilisticGraspOptimizer(hand, ee_link_names=doc["ee_link_name"], ee_link_offsets=doc["ee_link_offset"],
                                      palm_offset=palm, num_iters=iters, optimize_target=True, ref_q=doc["ref_q"], mass=0.1,
                                      com=[0.0, 0.0, 0.0], gravity=True, uncertainty=20.0, optimize_palm=True, device=dev,
                                      noise_seed=2024, env_offset=lo)
    lib = _lib.load()
    q0d, T0d, k0d = q0.to(dev), T0.to(dev), k0.to(dev)
    best = [None]

    def step():
        opt._init(q0d, T0d, k0d, 1, gp)
        opt.step(iters)
        best[0] = global_argmin(opt, world)

    for _ in range(args.warmup):
        step()
    l0 = lib.cdx_launch_count(gp._ctx)
    ms_step, clocks = timed(step, args.steps, world, rank, local)
    launches = lib.cdx_launch_count(gp._ctx) - l0
    nan_ci, nan_first = opt.nan_count()

    # e2e: the public call with HOST tensors (pinned): H2D of the candidates, D2H of the 5-tuple + arg-min
    hq, hT, hk = q0.pin_memory(), T0.pin_memory(), k0.pin_memory()

    def step_e2e():
        res = opt.optimize(hq, hT, hk, 1, gp, verbose=False)
        best[0] = global_argmin(opt, world)
        return [t.cpu() for t in res]

    step_e2e()
    ms_e2e, _ = timed(step_e2e, args.steps, world, rank, local, sample_clocks=False)

    # one more optimisation with CUDA events around every launch (cdx_pregrasp_profile): direct launches in the SAME candidate
    # groups and launch shapes as the timed loop (the timed loop replays them from a CUDA graph)
    _lib.check(lib.cdx_pregrasp_profile(gp._ctx, 1))
    piters = min(iters, 100)
    opt._init(q0d, T0d, k0d, 1, gp)
    opt.step(piters)
    kms = (ctypes.c_double * 4)()
    nl = ctypes.c_int()
    _lib.check(lib.cdx_pregrasp_profile_read(gp._ctx, kms, ctypes.byref(nl), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    _lib.check(lib.cdx_pregrasp_profile(gp._ctx, 0))
    # the extended query ALONE at the same size (one group's queries, nothing else on the GPU): what the kernel does when it does
    # not share the SMs with the other group's step / mean kernels
    groups_p = max(1, nl.value // piters)
    nq = (B + groups_p - 1) // groups_p * 4
    Xa = ((torch.rand(nq, 3, generator=torch.Generator().manual_seed(5), dtype=torch.float64) * 2 - 1) * 0.1).to(q0d.device)
    Xa += gp.X1.to(q0d.device).double().mean(dim=0)          # a 0.2 m box around the object
    o = [torch.empty(nq, dtype=torch.float64, device=q0d.device), torch.empty(nq, 3, dtype=torch.float64, device=q0d.device),
         torch.empty(nq, dtype=torch.float64, device=q0d.device), torch.empty(nq, 3, dtype=torch.float64, device=q0d.device)]
    st_ = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def ext_alone():
        _lib.check(lib.cdx_gpis_query(gp._ctx, Xa.data_ptr(), nq, o[0].data_ptr(), o[1].data_ptr(), o[2].data_ptr(), o[3].data_ptr(), st_))
    for _ in range(3):
        ext_alone()
    ea0, ea1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ea0.record()
    for _ in range(20):
        ext_alone()
    ea1.record()
    torch.cuda.synchronize()
    ext_alone_ms = ea0.elapsed_time(ea1) / 20
    if rank != 0:
        return None
    groups = max(1, nl.value // piters)
    per_launch = {k: kms[i] / max(1, nl.value) for i, k in enumerate(["wait_for_mean_query", "gpis_var_ext", "gpis_mean_overlapped", "pg_step"])}
    total_units = total * iters
    pipes = pipe_peaks()
    peaks, src = measured_peaks()
    ext_queries = (B + groups - 1) // groups * 4                  # per launch: one group's candidates x 4 contact points (U = 1)
    flop_ext = 2 * N * N + 34 * N                                  # extended query {mean, grad, std, grad std}
    achieved = flop_ext * ext_queries / (per_launch["gpis_var_ext"] * 1e-3) / 1e12
    peak = pipes.get("dmma_m8n8k4_tflops")
    crit = per_launch["gpis_var_ext"] + per_launch["wait_for_mean_query"] + per_launch["pg_step"]   # one group's critical path
    share = {k: per_launch[k] / crit for k in ("gpis_var_ext", "wait_for_mean_query", "pg_step")}
    ms_iter = ms_step / iters
    return {
        "metric": METRIC, "value": total_units / (ms_step * 1e-3), "unit": "candidate-iters/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(tag, hand, obj, N, args.candidates if weak else total, iters, world, weak),
                   "candidates_per_gpu": B, "ms_per_iteration": ms_iter,
                   "l2_policy": "latency-bound loop: the < 1.3 MB state (L^-1, E11^-1) and the per-candidate buffers are re-used "
                                "every iteration by design (resident working set), no flush applies",
                   "argmin": {"value": best[0][0], "global_index": best[0][1]},
                   "nan_candidate_iterations": nan_ci, "nan_first_iteration": nan_first,
                   "loop": "CUDA graph of 50-iteration chunks, %d candidate group(s) on separate stream pairs" % groups},
        "e2e": {"value": total_units / (ms_e2e * 1e-3), "unit": "candidate-iters/s",
                "h2d_bytes_per_step": int(B * (16 + 12 + 4 + 6) * 8), "d2h_bytes_per_step": int(B * (16 + 4 + 12 + 6 + 4) * 8 + 16)},
        "gpu_launches": int(launches), "clocks": clocks,
        "roofline": {"bound": "tensor", "kernel": "gpis_extw_kernel<TPS> (FP64 DMMA; warp-specialised extended query) inside the pregrasp "
                                                   "iteration, timed beside the other candidate group's kernels",
                     "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": (achieved / peak) if peak else None,
                     "peak_source": "FP64 DMMA m8n8k4 rate measured in this run by csrc/cdx_pipebench (MEASURED_PEAKS.json has no "
                                    "FP64 figure; its bf16 %s = %.0f TF/s)" % (src, peaks.get("bf16_tflops", 0)),
                     "flop_per_query": flop_ext, "queries_per_launch": ext_queries, "launch_groups": groups,
                     "traffic": ncu_traffic(args.ncu_summary_pg, "gpis_extw_kernel"),
                     "kernel_alone": {"ms": ext_alone_ms, "queries": nq,
                                      "achieved": flop_ext * nq / (ext_alone_ms * 1e-3) / 1e12,
                                      "frac": (flop_ext * nq / (ext_alone_ms * 1e-3) / 1e12 / peak) if peak else None,
                                      "note": "same launch size, nothing else on the GPU"},
                     "kernel_ms_per_launch": per_launch, "kernel_share_of_group_critical_path": share,
                     "iteration_ms_vs_sum_of_group_critical_paths": {"timed_loop": ms_iter, "sum": crit * groups},
                     "note": "per-launch times from %d extra iterations with CUDA events around each launch, in the timed loop's own "
                             "launch shape (%d group(s) of %d candidates, each walking its iterations on its own stream pair, so a "
                             "group's launches overlap the other's); pg_step_kernel is latency-bound scalar FP64 (16 lanes per "
                             "candidate)" % (piters, groups, (B + groups - 1) // groups)},
    }


# ------------------------------------------------------------------------------------------ CPU side (reference arm)
def cpu_pregrasp(hand, obj, total, B, iters, threads, kind):
    """The reference's own ProbabilisticGraspOptimizer.optimize (kind "reference", through oracle/ref_harness.py) or the oracle
    port on the host cores: the first B candidates of the workload's batch for `iters` iterations.
    -> candidate-iters/s, seconds"""
    from compliancedex_b200 import hand_path, state_path
    torch.set_num_threads(threads)
    doc, q0, T0, k0, palm, N = _candidates(hand, obj, total)
    q0, T0, k0, palm = q0[:B].clone(), T0[:B].clone(), k0[:B].clone(), palm[:B].copy()
    if kind == "reference":
        import contextlib
        import io
        from oracle import ref_harness as rh
        g = rh.load_gpis(f"{obj}_state")
        ropt, _ = rh.make_prob_optimizer(hand, palm, iters, mass=0.1, com=[0.0, 0.0, 0.0])
        t0 = time.perf_counter()
        with contextlib.redirect_stdout(io.StringIO()), rh.pinned_kabsch_noise(2024):
            ropt.optimize(q0, T0, k0, 1, g)
        dt = time.perf_counter() - t0
    else:
        from oracle.gpis_oracle import GPISOracle
        from oracle.pregrasp_oracle import PregraspOracle
        go = GPISOracle(0.08, 1.0).load_npz(state_path(f"{obj}_state"))
        o = PregraspOracle(hand_path(hand), palm, iters, mass=0.1, noise_seed=2024)
        t0 = time.perf_counter()
        o.optimize(q0, T0, k0, 1, go)
        dt = time.perf_counter() - t0
    return B * iters / dt, dt, N


def ref_pregrasp(args, threads, world, tag):
    import warnings
    warnings.filterwarnings("ignore")
    kind = reference_kind()
    weak = tag == "pregrasp_c2"
    if weak:
        hand, obj, iters, total, Bs, its = args.hand, args.object, args.iters, args.candidates, 64, 10
    else:
        hand, obj, iters, total, Bs, its = "allegro", args.c5_object, args.c5_iters, args.c5_candidates, 64, 6
    vals = []
    for i in range(args.warmup + args.steps):
        v, dt, N = cpu_pregrasp(hand, obj, total, Bs, its, threads, kind)
        if i >= args.warmup:
            vals.append((v, dt))
    v = float(np.mean([a for a, _ in vals]))
    ms = float(np.mean([b for _, b in vals])) * 1e3
    sample = ("each step = the first %d candidates x %d iterations of the same batch through %s on the host cores (%.1f s)"
              % (Bs, its, "the reference's own optimize_pregrasp.py / gpis.py / FK package" if kind == "reference"
                 else "the oracle port of the reference (torch CPU)", ms * 1e-3))
    return {"impl": "reference", "metric": METRIC, "value": v, "unit": "candidate-iters/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak" if weak else "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(tag, hand, obj, N, total, iters, world, weak), "sample": sample},
            "cpu_baseline": {"value": v, "unit": "candidate-iters/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "candidate-iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
