"""bench.py's GPIS workloads (BASELINE.json configs[0], [2], [3]):
  gpis_c3 : synthetic TPS GPIS N=4096, 10 M unit queries {mean, grad_mean, std} per step, sharded over ranks (strong scaling)
  gpis_c1 : banana state (N=361), the 64^3 grid of GPIS.get_visualization_data
  fit_c4  : GPIS.fit (assembly + blocked Cholesky + inverses) at N=16,384
GPU legs go through the C ABI / the drop-in class; the CPU legs run the reference's own gpis.py (kind "reference") when it
is available (oracle/ref_harness.py), else the oracle port."""
import ctypes
import os
import time

import numpy as np
import torch

from bench_common import (ClockSampler, barrier, i8_peak, max_over_ranks, measured_peaks, ncu_traffic, pipe_peaks, reference_kind,
                          timed)

ROOT = os.path.dirname(os.path.abspath(__file__))
METRIC_Q = "GPIS mean+var+grad queries/sec"


def _wl_c3(N, M):
    return ("gpis_c3: synthetic TPS GPIS N=%d, %d unit queries {mean,grad_mean,std} per step, sharded by contiguous ranges" % (N, M))


def _wl_c1():
    return "gpis_c1: banana_state GPIS (N=361), get_visualization_data over the 64^3 grid of its +-0.15 box (262,144 queries)"


def _wl_fit(N):
    return "fit_c4: GPIS.fit on N=%d synthetic observed+completed points (train_gpis_completion.py shape), 1 GPU" % N


# ------------------------------------------------------------------------------------------ gpis_c3 (GPU)
def run_gpis_c3(args, rank, world, local):
    from compliancedex_b200 import GPIS, _lib
    from compliancedex_b200.distributed import shard_range
    from compliancedex_b200.synthetic import synthetic_queries, synthetic_training_set
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    N, total_q = args.ntrain, args.queries
    X1, y1, noise = synthetic_training_set(N, seed=0)
    lo, hi = shard_range(total_q, rank, world)
    Xh = synthetic_queries(total_q, lo, hi)
    M = Xh.shape[0]
    g = GPIS(0.08, 1.0)
    g.set_var_mode(args.var_mode)
    X1d, y1d, nzd = X1.to(dev), y1.to(dev), noise.to(dev)
    g.fit(X1d, y1d, nzd)                                   # warm (module load, attribute setup)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    g.fit(X1d, y1d, nzd)
    torch.cuda.synchronize()
    fit_s = time.perf_counter() - t0
    lib = _lib.load()
    Xd = Xh.to(dev)
    mean = torch.empty(M, dtype=torch.float64, device=dev)
    std = torch.empty_like(mean)
    gm = torch.empty(M, 3, dtype=torch.float64, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def step():
        _lib.check(lib.cdx_gpis_query(g._ctx, Xd.data_ptr(), M, mean.data_ptr(), gm.data_ptr(), std.data_ptr(), None, st))

    for _ in range(args.warmup):
        step()
    l0 = lib.cdx_launch_count(g._ctx)
    ms_step, clocks = timed(step, args.steps, world, rank, local)
    launches = lib.cdx_launch_count(g._ctx) - l0

    # e2e: host buffers through the C ABI (H2D of the queries + D2H of all outputs inside the timed region)
    Xp = Xh.contiguous().pin_memory()
    mh = torch.empty(M, dtype=torch.float64).pin_memory()
    sh = torch.empty(M, dtype=torch.float64).pin_memory()
    gh = torch.empty(M, 3, dtype=torch.float64).pin_memory()

    def step_host():
        _lib.check(lib.cdx_gpis_query_host(g._ctx, Xp.data_ptr(), M, mh.data_ptr(), gh.data_ptr(), sh.data_ptr(), None, st))

    step_host()
    ms_e2e, _ = timed(step_host, args.steps, world, rank, local, sample_clocks=False)

    # dominant kernel alone (std only -> one launch), CUDA events on the launching stream
    def step_std():
        _lib.check(lib.cdx_gpis_query(g._ctx, Xd.data_ptr(), M, None, None, std.data_ptr(), None, st))

    step_std()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step_std()
    e1.record()
    torch.cuda.synchronize()
    kernel_ms = e0.elapsed_time(e1) / args.steps
    if rank != 0:
        return None
    Np = (N + 63) // 64 * 64
    flop_per_query = N * N + 26 * N                                # SURVEY 8(d) unit query
    pipes = pipe_peaks()
    peaks, src = measured_peaks()
    fp64_equiv = flop_per_query * M / (kernel_ms * 1e-3) / 1e12    # algorithmic FP64 flop of this rank's launch / its time
    dmma_peak = pipes.get("dmma_m8n8k4_tflops")
    i8 = args.var_mode != -1 and 1024 <= Np <= 8192
    if i8:
        S = args.var_mode if args.var_mode >= 4 else 5
        # INT8 multiply-adds the kernel ISSUES per query (cdx_query_i8x.cu, CTA pair): S(S+1)/2 slice pairs x 128-row tiles x
        # 32-point stages; round r of a pair walks the k chunks up to the diagonal of its LAST tile
        NT, NC = (Np + 127) // 128, Np // 32
        stages = sum(min((r * 2 + 2) * 4, NC) for r in range((NT + 1) // 2)) * 2
        ops_per_query = 2 * (S * (S + 1) // 2) * 128 * 32 * stages
        useful_ops_per_query = 2 * (S * (S + 1) // 2) * (N * (N + 1) // 2)
        achieved = ops_per_query * M / (kernel_ms * 1e-3) / 1e12
        peak_burst = i8_peak(pipes)
        peak = i8_peak(pipes, sustained=True) or peak_burst     # the std-only step below is ~1 s of back-to-back tensor work
        roof = {"bound": "tensor", "kernel": "gpis_std_i8x_kernel<%d, CTA pair> (tcgen05.mma.cta_group::2 kind::i8, INT32 accumulators in TMEM); "
                                           "its time here includes i8x_kstar_kernel, which generates the k* byte slices once per query chunk" % S,
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
                "peak_source": "tcgen05.mma kind::i8 issue rate MEASURED in this run by csrc/cdx_pipebench, SUSTAINED figure (M128 x N256 "
                               "x K32 from shared memory, back to back for ~3 s, second half timed: the kernel is timed inside a long "
                               "step under the power cap; MEASURED_PEAKS.json has no INT8 figure - its bf16 %s = %.0f TF/s burst, %.0f "
                               "sustained); achieved counts the INT8 multiply-adds the kernel issues x 2 (TOP/s)"
                               % (src, peaks.get("bf16_tflops", 0), peaks.get("bf16_tflops_sustained", 0)),
                "peak_burst": peak_burst, "frac_of_burst_peak": achieved / peak_burst if peak_burst else None,
                "int8_ops_per_query": ops_per_query, "int8_ops_per_query_lower_triangle_only": useful_ops_per_query, "slices": S,
                "algorithmic_fp64_tflops": fp64_equiv, "fp64_dmma_peak_tflops": dmma_peak,
                "speedup_over_fp64_dmma_peak_rate": fp64_equiv / dmma_peak if dmma_peak else None,
                "note": "the FP64 contraction L^-1 k* (N^2 flop/query) is evaluated as error-free INT8 slice products "
                        "(S(S+1)/2 byte-slice pairs); FP64 instructions and tcgen05 MMAs exclude each other on this chip "
                        "(csrc/cdx_mixbench.cu, DESIGN.md section 4), so the k* slices are generated ONCE per query by a separate "
                        "FP64 kernel into the B-operand image and the MMA kernel is pure INT8; mean / grad_mean are "
                        "accumulated by that same k* kernel (same distances) and reduced by i8x_mean_reduce_kernel"}
    else:
        roof = {"bound": "tensor", "kernel": "gpis_var_kernel<TPS,128,unit> (FP64 DMMA)", "achieved": fp64_equiv,
                "peak": dmma_peak, "unit": "TFLOP/s", "frac": (fp64_equiv / dmma_peak) if dmma_peak else None,
                "peak_source": "FP64 DMMA m8n8k4 rate measured in this run by csrc/cdx_pipebench (MEASURED_PEAKS.json "
                               "has no FP64 figure; its bf16 %s = %.0f TF/s)" % (src, peaks.get("bf16_tflops", 0))}
    roof.update({"flop_per_query": flop_per_query, "queries_per_launch": M, "kernel_ms": kernel_ms,
                 "kernel_share_of_step": kernel_ms / ms_step,
                 "traffic": ncu_traffic(args.ncu_summary_c3, "gpis_std_i8") if i8 else None, "pipes": pipes})
    e2e_s = ms_e2e * 1e-3
    return {
        "metric": METRIC_Q, "value": total_q / (ms_step * 1e-3), "unit": "queries/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": _wl_c3(N, total_q), "var_mode": args.var_mode,
                   "l2_policy": "inputs+outputs (%.0f MB per rank) and the sliced L^-1 operand (%.0f MB) exceed L2; no flush needed"
                                % (M * 64 / 1e6, (5 * Np * Np if i8 else 8 * N * N) / 1e6),
                   "fit_seconds": fit_s},
        "e2e": {"value": total_q / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": int(M * 24), "d2h_bytes_per_step": int(M * 40),
                "value_including_one_fit": total_q / (e2e_s + fit_s)},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roof,
    }


# ------------------------------------------------------------------------------------------ gpis_c1 (GPU)
def run_gpis_c1(args, rank, world, local):
    from compliancedex_b200 import GPIS, _lib, state_path
    from oracle.gpis_oracle import viz_grid   # grid layout only (test/bench helper; no GPIS arithmetic)
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    lb, ub, steps = [-0.15] * 3, [0.15] * 3, 64
    g = GPIS(0.08, 1.0)
    g.load_state_data(state_path("banana_state"), device=dev)
    N = int(g.X1.shape[0])
    lib = _lib.load()
    X = viz_grid(lb, ub, steps).reshape(-1, 3).to(dev).contiguous()
    M = X.shape[0]
    mean = torch.empty(M, dtype=torch.float64, device=dev)
    std = torch.empty_like(mean)
    gm = torch.empty(M, 3, dtype=torch.float64, device=dev)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2

    def step():
        flush.zero_()                                                  # the 6 MB grid + 10 MB of outputs would otherwise stay in L2
        _lib.check(lib.cdx_gpis_query(g._ctx, X.data_ptr(), M, mean.data_ptr(), gm.data_ptr(), std.data_ptr(), None, st))

    def step_flush_only():
        flush.zero_()

    for _ in range(args.warmup):
        step()
    l0 = lib.cdx_launch_count(g._ctx)
    ms_all, clocks = timed(step, args.steps, world, rank, local)
    launches = lib.cdx_launch_count(g._ctx) - l0
    ms_flush, _ = timed(step_flush_only, args.steps, world, rank, local, sample_clocks=False)
    ms_step = max(ms_all - ms_flush, 1e-6)

    def step_e2e():                                     # the public call: axis vectors up, grid expanded on the device, queried,
                                                        # float32 planes read back through one pinned buffer
        return g.get_visualization_data(lb, ub, steps)

    step_e2e()
    ms_e2e, _ = timed(step_e2e, args.steps, world, rank, local, sample_clocks=False)
    if rank != 0:
        return None
    pipes = pipe_peaks()
    flop = N * N + 26 * N
    achieved = flop * M / (ms_step * 1e-3) / 1e12
    peak = pipes.get("dmma_m8n8k4_tflops")
    return {"metric": METRIC_Q, "value": M / (ms_step * 1e-3), "unit": "queries/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "shipped state banana_state.npz, synthetic grid",
            "config": {"workload": _wl_c1(), "l2_policy": "256 MB flush buffer written before every step; its time (%.3f ms) "
                                                            "measured separately and subtracted" % ms_flush},
            "e2e": {"value": M / (ms_e2e * 1e-3), "unit": "queries/s", "h2d_bytes_per_step": int(3 * steps * 4),
                    "d2h_bytes_per_step": int(M * 5 * 4)},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "gpis_var_kernel<TPS,128,unit,128> (FP64 DMMA)", "achieved": achieved,
                         "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
                         "peak_source": "FP64 DMMA m8n8k4 rate measured in this run by csrc/cdx_pipebench",
                         "flop_per_query": flop, "queries_per_launch": M, "kernel_ms": ms_step,
                         "traffic": ncu_traffic("r02_ncu_unit_c1_summary.json", "gpis_var_kernel")}}


# ------------------------------------------------------------------------------------------ fit_c4 (GPU)
def run_fit_c4(args, rank, world, local):
    from compliancedex_b200 import GPIS, _lib
    from compliancedex_b200.synthetic import synthetic_training_set
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    N = args.fit_n
    X1, y1, noise = synthetic_training_set(N, seed=0)
    X1d, y1d, nzd = X1.to(dev), y1.reshape(-1).contiguous().to(dev), noise.to(dev)
    lib = _lib.load()
    ctx = _lib.new_context(local)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def step():
        _lib.check(lib.cdx_gpis_fit(ctx, X1d.data_ptr(), y1d.data_ptr(), nzd.data_ptr(), 0.0, N, 1.0, 0, 0.08, st))

    for _ in range(args.warmup):
        step()
    _lib.check(lib.cdx_gpis_fit_profile(ctx, 1))
    l0 = lib.cdx_launch_count(ctx)
    ms_step, clocks = timed(step, args.steps, world, rank, local)
    launches = lib.cdx_launch_count(ctx) - l0
    ph = (ctypes.c_double * 5)()
    _lib.check(lib.cdx_gpis_fit_profile_read(ctx, ph, st))
    _lib.check(lib.cdx_gpis_fit_profile(ctx, 0))
    ctx.close()

    # e2e: the public GPIS.fit from pinned HOST tensors: H2D of the training set, fit, E11 attribute, R read back
    Xp, yp, nzp = X1.pin_memory(), y1.pin_memory(), noise.pin_memory()
    g = GPIS(0.08, 1.0)

    def step_e2e():
        g.fit(Xp.to(dev, non_blocking=True), yp.to(dev, non_blocking=True), nzp.to(dev, non_blocking=True))
        return float(g.R)

    step_e2e()
    ms_e2e, _ = timed(step_e2e, args.steps, world, rank, local, sample_clocks=False)
    if rank != 0:
        return None
    pipes = pipe_peaks()
    peak = pipes.get("dmma_m8n8k4_tflops")
    names = ["assembly", "cholesky", "inverse_factor", "inverse_matrix", "alpha_and_pack"]
    phase_ms = {n: ph[i] for i, n in enumerate(names)}
    third = N ** 3 / 3.0
    tf = lambda ms: third / (ms * 1e-3) / 1e12 if ms > 0 else None   # noqa: E731
    phase_tf = {n: tf(phase_ms[n]) for n in ("cholesky", "inverse_factor", "inverse_matrix")}
    achieved = phase_tf["cholesky"]
    return {"metric": "GPIS fits/sec (N=%d)" % N, "value": 1.0 / (ms_step * 1e-3), "unit": "fits/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": _wl_fit(N), "l2_policy": "the %.1f GB matrix exceeds L2; no flush needed" % (8.0 * N * N / 1e9)},
            "e2e": {"value": 1.0 / (ms_e2e * 1e-3), "unit": "fits/s", "h2d_bytes_per_step": int(N * 5 * 8), "d2h_bytes_per_step": 12,
                    "ms_per_fit": ms_e2e},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": "cdx_dgemm (FP64 DMMA) in the blocked Cholesky phase", "achieved": achieved,
                         "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if (peak and achieved) else None,
                         "peak_source": "FP64 DMMA m8n8k4 rate measured in this run by csrc/cdx_pipebench",
                         "flop_per_phase": third, "phase_ms": phase_ms, "phase_tflops": phase_tf,
                         "phase_frac_of_dmma_peak": {k: (v / peak if (v and peak) else None) for k, v in phase_tf.items()},
                         "whole_fit_tflops_over_N3": 3 * third / (ms_step * 1e-3) / 1e12,
                         "traffic": ncu_traffic("r02_ncu_fit16k_summary.json", "trailing update")}}


# ------------------------------------------------------------------------------------------ CPU legs (reference arm)
def _ref_gpis_object(kind):
    if kind == "reference":
        from oracle import ref_harness as rh
        rh.install()
        from gpis import GPIS
        return GPIS(0.08, 1.0)
    from oracle.gpis_oracle import GPISOracle
    return GPISOracle(0.08, 1.0)


def _ref_gpis_state(kind, name):
    if kind == "reference":
        from oracle import ref_harness as rh
        return rh.load_gpis(f"{name}_state")
    from compliancedex_b200 import state_path
    from oracle.gpis_oracle import GPISOracle
    return GPISOracle(0.08, 1.0).load_npz(state_path(f"{name}_state"))


def _line(impl_metric, unit, value, ms, workload, sample, kind, threads, args, world, extra=None, dtype="f64"):
    cb = {"value": value, "unit": unit, "cores": threads, "kind": kind, "sample": sample}
    if extra:
        cb.update(extra)
    return {"impl": "reference", "metric": impl_metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": dtype, "data": "synthetic", "config": {"workload": workload, "sample": sample}, "cpu_baseline": cb,
            "e2e": {"value": value, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def ref_gpis_c3(args, threads, world):
    """gpis.py on the host cores: fit (N=4096) once, then each step = pred + compute_normal on a `sample`-query chunk - the
    largest call the reference can make is a few thousand queries (it materialises the MxM covariance, gpis.py:57-58), and
    every call refactorises (LU at :53, inverse at :77)."""
    from oracle.gpis_oracle import synthetic_training_set
    kind = reference_kind()
    torch.set_num_threads(threads)
    N, sample = args.ntrain, 2048
    X1, y1, noise = synthetic_training_set(N, seed=0)
    o = _ref_gpis_object(kind)
    t0 = time.perf_counter()
    o.fit(X1, y1, noise)
    fit_s = time.perf_counter() - t0
    X = (torch.rand(sample, 3, generator=torch.Generator().manual_seed(1234), dtype=torch.float64) * 2 - 1) * 0.1
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        o.pred(X)
        o.compute_normal(X)
        if i >= args.warmup:
            times.append(time.perf_counter() - t0)
    dt = float(np.mean(times))
    t0 = time.perf_counter()                                   # the per-call factorisation share: the same pair of calls on 16 queries
    o.pred(X[:16])
    o.compute_normal(X[:16])
    dt_small = time.perf_counter() - t0
    v = sample / dt
    smp = ("each step = gpis.py pred (LU of the %d^2 matrix + MxM covariance) + compute_normal (explicit inverse) on %d queries; "
           "%.2f s per step, of which %.2f s is the per-call refactorisation (same calls on 16 queries); marginal rate "
           "%.3g queries/s" % (N, sample, dt, dt_small, (sample - 16) / max(dt - dt_small, 1e-9)))
    return _line(METRIC_Q, "queries/s", v, dt * 1e3, _wl_c3(N, args.queries), smp, kind, threads, args, world,
                 {"fit_seconds": fit_s, "per_call_refactorisation_seconds": dt_small,
                  "marginal_queries_per_s": (sample - 16) / max(dt - dt_small, 1e-9)})


def ref_gpis_c1(args, threads, world):
    kind = reference_kind()
    torch.set_num_threads(threads)
    o = _ref_gpis_state(kind, "banana")
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        o.get_visualization_data([-0.15] * 3, [0.15] * 3, steps=64)
        if i >= args.warmup:
            times.append(time.perf_counter() - t0)
    dt = float(np.mean(times))
    smp = "the whole workload: get_visualization_data(steps=64) = 64 slabs x (pred + compute_normal) (gpis.py:113-125), %.1f s" % dt
    return _line(METRIC_Q, "queries/s", 64 ** 3 / dt, dt * 1e3, _wl_c1(), smp, kind, threads, args, world)


def ref_fit_c4(args, threads, world):
    """The reference's fit only assembles E11 (gpis.py:33-40); the factorisation cost it defers to every pred (LU, :53) is
    timed beside it, and torch.linalg.cholesky of the same matrix as the like-for-like factorisation."""
    from oracle.gpis_oracle import synthetic_training_set
    kind = reference_kind()
    torch.set_num_threads(threads)
    N = args.fit_n
    X1, y1, noise = synthetic_training_set(N, seed=0)
    times = []
    for i in range(args.warmup + args.steps):
        o = _ref_gpis_object(kind)
        t0 = time.perf_counter()
        o.fit(X1, y1, noise)
        t_asm = time.perf_counter() - t0
        t0 = time.perf_counter()
        torch.linalg.cholesky(o.E11)
        t_chol = time.perf_counter() - t0
        if i >= args.warmup:
            times.append((t_asm, t_chol))
    t_asm, t_chol = (float(np.mean([t[k] for t in times])) for k in (0, 1))
    Xq = (torch.rand(64, 3, generator=torch.Generator().manual_seed(5), dtype=torch.float64) * 2 - 1) * 0.1
    t0 = time.perf_counter()
    o.pred(Xq)
    t_pred = time.perf_counter() - t0
    dt = t_asm + t_chol
    smp = ("the whole workload once per step: gpis.py fit (assembly, %.1f s) + torch.linalg.cholesky(E11) (%.1f s); one gpis.py "
           "pred call on 64 queries (the LU the reference pays per call) takes %.1f s" % (t_asm, t_chol, t_pred))
    return _line("GPIS fits/sec (N=%d)" % N, "fits/s", 1.0 / dt, dt * 1e3, _wl_fit(N), smp, kind, threads, args, world,
                 {"assembly_seconds": t_asm, "cholesky_seconds": t_chol, "one_pred_lu_seconds": t_pred})
