#!/usr/bin/env python
"""bench.py - hot-path throughput of compliancedex_b200 on B200 (contract in the task brief).

Workloads (BASELINE.json configs):
  pregrasp_c2 : ProbabilisticGraspOptimizer, Leap hand on the mug GPIS, 1,024 candidates x 200 iterations  [configs[1]]
  gpis_c3     : synthetic GPIS N=4096, 10M unit queries {mean, grad_mean, std}, sharded over ranks   [configs[2]]
One JSON line on rank 0.  `--impl reference` times the CPU restatement of the reference's algorithm (oracle/, kind
"port": /root/reference cannot travel to the GPU box) on the host cores instead.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np
import torch


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cdx", choices=["cdx", "reference"])
    ap.add_argument("--workload", default=os.environ.get("CDX_BENCH_WORKLOAD", "auto"))
    ap.add_argument("--queries", type=int, default=10_000_000)
    ap.add_argument("--ntrain", type=int, default=4096)
    ap.add_argument("--candidates", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=200)
    ap.add_argument("--hand", default="leap", choices=["leap", "allegro"])
    ap.add_argument("--object", default="mug", help="shipped GPIS state of the pregrasp workload (mug, hammer, lego, ...)")
    ap.add_argument("--points", type=int, default=1_000_000, help="sdf workload: query points")
    ap.add_argument("--subdiv", type=int, default=5, help="sdf workload: icosphere subdivisions (20 * 4^n faces)")
    ap.add_argument("--sdf-mode", type=int, default=0, help="sdf workload: 0 auto (culled search), 1 exhaustive scan, 2 culled")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--var-mode", type=int, default=0,
                    help="gpis_c3: 0 auto (tcgen05 INT8 slices), -1 FP64 tensor MMA, 4..6 INT8 with that many slices")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------ helpers
class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region: NVML in a thread every 5 ms (the pregrasp workload's
    timed region is tens of milliseconds - an `nvidia-smi -lms` child process would not even have started), falling back to
    an nvidia-smi loop when pynvml is unusable."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.nv, self.h = index, [], None, None, None
        self.sm, self.mx, self.bits, self.stop_flag = [], None, 0, False
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                self.h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode() if not uuid.startswith("GPU-") else uuid.encode())
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
                self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.nv = pynvml
        except Exception:
            self.nv = None

    def _poll(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            time.sleep(0.005)

    def start(self):
        if self.nv is not None:
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nv is not None:
            self.stop_flag = True
            self.thread.join(timeout=2)
            nv = self.nv
            masks = [("hw_slowdown", getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8)),
                     ("hw_thermal_slowdown", getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40)),
                     ("sw_thermal_slowdown", getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20)),
                     ("sw_power_cap", getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4))]
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                    "reasons": [n for n, m in masks if self.bits & m], "samples": len(self.sm), "source": "nvml, 5 ms period"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": "nvidia-smi -lms 100"}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


def ncu_traffic(summary_file, kernel_substr):
    """DRAM bytes (read + write) per launch of a kernel from the committed ncu --set full summary under profiles/
    (scripts/ncu_summary.py); None when no capture of that kernel is committed."""
    p = os.path.join(ROOT, "profiles", summary_file)
    if not os.path.isfile(p):
        return None
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for key, d in json.load(open(p)).items():
        if kernel_substr in key:
            tot = 0.0
            for f in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                v, u = d.get(f, "0 byte").split()[:2]
                tot += float(v) * mult.get(u, 1.0)
            return {"bytes_per_launch": tot, "source": "profiles/%s (%s, grid %s)" % (summary_file, key.split(":")[0],
                                                                                     d.get("launch__grid_size", "?").strip())}
    return None


def fp64_pipe_peaks():
    """FP64 pipe rates are not in MEASURED_PEAKS.json; measure them here with the repo's own microbenchmark."""
    exe = os.path.join(ROOT, "compliancedex_b200", "csrc", "build", "cdx_pipebench")
    try:
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120).stdout.strip().splitlines()[-1]
        return json.loads(out)
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)}


def dist_setup(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        backend = "nccl" if args.impl == "cdx" else "gloo"
        if args.impl == "cdx":
            torch.cuda.set_device(local)
        # NCCL prints its version banner to STDOUT when the first communicator is created: send fd 1 to stderr while the process
        # group comes up (init + one warm-up collective), so that stdout carries exactly one JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group(backend=backend, device_id=torch.device("cuda", local) if backend == "nccl" else None)
            if backend == "nccl":
                t = torch.zeros(1, device="cuda")
                dist.all_reduce(t)
                torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    return rank, world, local


def max_over_ranks(ms, world):
    if world == 1:
        return ms
    import torch.distributed as dist
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


# ------------------------------------------------------------------------------------------ GPIS C3
def gpis_c3_inputs(ntrain, m_total, rank, world):
    from compliancedex_b200.distributed import shard_range
    from compliancedex_b200.synthetic import synthetic_queries, synthetic_training_set
    X1, y1, noise = synthetic_training_set(ntrain, seed=0)
    lo, hi = shard_range(m_total, rank, world)
    return X1, y1, noise, synthetic_queries(m_total, lo, hi), hi - lo


def run_gpis_c3(args, rank, world, local):
    import ctypes
    from compliancedex_b200 import GPIS, _lib
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    X1, y1, noise, Xh, per = gpis_c3_inputs(args.ntrain, args.queries, rank, world)
    M = Xh.shape[0]
    g = GPIS(0.08, 1.0)
    g.set_var_mode(args.var_mode)
    t0 = time.perf_counter()
    g.fit(X1.to(dev), y1.to(dev), noise.to(dev))
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
    sampler = ClockSampler(local)
    l0 = lib.cdx_launch_count(g._ctx)
    barrier(world)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier(world)
    ms = max_over_ranks(e0.elapsed_time(e1), world)
    clocks = sampler.stop() if rank == 0 else None
    launches = lib.cdx_launch_count(g._ctx) - l0

    # e2e: host buffers through the C ABI (H2D of the queries + D2H of all outputs inside the timed region)
    Xp = Xh.contiguous().pin_memory()
    mh = torch.empty(M, dtype=torch.float64).pin_memory()
    sh = torch.empty(M, dtype=torch.float64).pin_memory()
    gh = torch.empty(M, 3, dtype=torch.float64).pin_memory()

    def step_host():
        _lib.check(lib.cdx_gpis_query_host(g._ctx, Xp.data_ptr(), M, mh.data_ptr(), gh.data_ptr(), sh.data_ptr(), None, st))

    step_host()
    barrier(world)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        step_host()
    e1.record()
    barrier(world)
    ms_e2e = max_over_ranks(e0.elapsed_time(e1), world)

    # dominant kernel alone (std only -> one launch), CUDA events on the launching stream
    def step_std():
        _lib.check(lib.cdx_gpis_query(g._ctx, Xd.data_ptr(), M, None, None, std.data_ptr(), None, st))

    step_std()
    e0.record()
    for _ in range(args.steps):
        step_std()
    e1.record()
    torch.cuda.synchronize()
    kernel_ms = e0.elapsed_time(e1) / args.steps

    if rank != 0:
        return None
    N = args.ntrain
    Np = (N + 63) // 64 * 64
    total_q = args.queries
    ms_step = ms / args.steps
    flop_per_query = N * N + 26 * N                                # SURVEY 8(d) unit query
    pipes = fp64_pipe_peaks()
    peaks, src = measured_peaks()
    fp64_equiv = flop_per_query * M / (kernel_ms * 1e-3) / 1e12    # algorithmic FP64 flop of this rank's launch / its time
    dmma_peak = pipes.get("dmma_m8n8k4_tflops")
    i8 = args.var_mode != -1 and 1024 <= Np <= 8192
    if i8:
        S = args.var_mode if args.var_mode >= 4 else 5
        NT = Np // 64
        # executed INT8 multiply-adds: S(S+1)/2 slice pairs x 64x64 blocks of the lower triangle (diagonal blocks whole)
        ops_per_query = 2 * (S * (S + 1) // 2) * 64 * 64 * (NT * (NT + 1) // 2)
        achieved = ops_per_query * M / (kernel_ms * 1e-3) / 1e12
        peak = 2.0 * peaks.get("bf16_tflops", 0)
        roof = {"bound": "tensor", "kernel": "gpis_std_i8_kernel<%d> (tcgen05 kind::i8, INT32 accumulators in TMEM)" % S,
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
                "peak_source": "INT8 dense tensor peak taken as 2 x the %s bf16 figure of MEASURED_PEAKS.json (%.0f TF/s); "
                               "achieved counts executed INT8 multiply-adds x 2 (TOP/s)" % (src, peaks.get("bf16_tflops", 0)),
                "int8_ops_per_query": ops_per_query, "slices": S,
                "algorithmic_fp64_tflops": fp64_equiv, "fp64_dmma_peak_tflops": dmma_peak,
                "algorithmic_vs_fp64_dmma_peak": fp64_equiv / dmma_peak if dmma_peak else None,
                "note": "the FP64 contraction L^-1 k* (N^2 flop/query) is evaluated as error-free INT8 slice products; the "
                        "kernel is bound by the FP64 generation of the k* slices (16 warps, DFMA + MUFU.RSQ + PRMT), not by "
                        "the tensor pipe; mean / grad_mean come from the separate TMA-staged mean kernel"}
    else:
        roof = {"bound": "tensor", "kernel": "gpis_var_kernel<TPS,128,unit> (FP64 DMMA)", "achieved": fp64_equiv,
                "peak": dmma_peak, "unit": "TFLOP/s", "frac": (fp64_equiv / dmma_peak) if dmma_peak else None,
                "peak_source": "FP64 DMMA m8n8k4 rate measured in this run by csrc/cdx_pipebench (MEASURED_PEAKS.json "
                               "has no FP64 figure; its bf16 %s = %.0f TF/s)" % (src, peaks.get("bf16_tflops", 0))}
    roof.update({"flop_per_query": flop_per_query, "queries_per_launch": M, "kernel_ms": kernel_ms,
                 "kernel_share_of_step": kernel_ms / ms_step,
                 "traffic": ncu_traffic("r01_ncu_gpis_c3_i8_summary.json", "gpis_std_i8_kernel") if i8 else None,
                 "fp64_pipes": pipes})
    out = {
        "metric": "GPIS mean+var+grad queries/sec", "value": total_q / (ms_step * 1e-3), "unit": "queries/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "gpis_c3: synthetic TPS GPIS N=%d, %d unit queries {mean,grad_mean,std} per step, "
                               "sharded by contiguous ranges" % (N, total_q),
                   "var_mode": args.var_mode,
                   "l2_policy": "inputs+outputs (%.0f MB) and the L^-1 operand (%.0f MB) exceed L2; no flush needed"
                                % (M * 64 / 1e6, (5 * Np * Np if i8 else 8 * N * N) / 1e6),
                   "fit_seconds": fit_s},
        "e2e": {"value": total_q / (ms_e2e / args.steps * 1e-3), "unit": "queries/s",
                "h2d_bytes_per_step": int(M * 24), "d2h_bytes_per_step": int(M * 40)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof,
    }
    return out


def cpu_gpis_c3(ntrain, sample, threads):
    """Reference algorithm (oracle port): pred + compute_normal per chunk of `sample` queries (each call refactorises)."""
    from oracle.gpis_oracle import GPISOracle, synthetic_training_set
    torch.set_num_threads(threads)
    X1, y1, noise = synthetic_training_set(ntrain, seed=0)
    o = GPISOracle(0.08, 1.0).fit(X1, y1, noise)
    X = (torch.rand(sample, 3, generator=torch.Generator().manual_seed(1234), dtype=torch.float64) * 2 - 1) * 0.1
    t0 = time.perf_counter()
    o.pred(X)
    o.compute_normal(X)
    dt = time.perf_counter() - t0
    return sample / dt, dt


# ------------------------------------------------------------------------------------------ main
def main():
    args = parse()
    rank, world, local = dist_setup(args)
    workload = args.workload
    if workload == "auto":
        try:
            from bench_pregrasp import run_bench_c2  # noqa: F401
            workload = "pregrasp_c2"
        except ImportError:
            workload = "gpis_c3"
    threads = os.cpu_count() or 1

    if args.impl == "reference":
        if rank != 0:
            return
        if workload == "gpis_c3":
            vals = []
            for i in range(args.warmup + args.steps):
                v, dt = cpu_gpis_c3(args.ntrain, 2048, threads)
                if i >= args.warmup:
                    vals.append((v, dt))
            v = float(np.mean([a for a, _ in vals]))
            ms = float(np.mean([b for _, b in vals])) * 1e3
            line = {"impl": "reference", "metric": "GPIS mean+var+grad queries/sec", "value": v, "unit": "queries/s",
                    "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
                    "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                    "config": {"workload": "gpis_c3: synthetic TPS GPIS N=%d; each step = pred + compute_normal on a "
                                           "2048-query sample (the reference refactorises per call)" % args.ntrain},
                    "cpu_baseline": {"value": v, "unit": "queries/s", "cores": threads, "kind": "port",
                                     "sample": "2048 queries per step: one pred (LU) + one compute_normal (inverse)"},
                    "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        elif workload == "sdf":
            from bench_sdf import METRIC, cpu_sdf
            vals = [cpu_sdf(1000, args.subdiv, 1) for _ in range(args.warmup + args.steps)][args.warmup:]
            v, ms = float(np.mean([a for a, _ in vals])), float(np.mean([b for _, b in vals])) * 1e3
            line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
                    "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                    "dtype": "f32", "data": "synthetic",
                    "config": {"workload": "sdf: 1000-point sample against the icosphere of %d faces; the reference has no CPU "
                                           "implementation of this op, so this is the numpy port" % (20 * 4 ** args.subdiv)},
                    "cpu_baseline": {"value": v, "unit": "queries/s", "cores": 1, "kind": "port", "sample": "1000 points per step"},
                    "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        else:
            from bench_pregrasp import run_reference_c2
            line = run_reference_c2(args, threads, world)
        print(json.dumps(line), flush=True)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the cdx path has no CPU fallback")
    if workload == "gpis_c3":
        line = run_gpis_c3(args, rank, world, local)
        if rank == 0 and not args.no_cpu_baseline:
            v, dt = cpu_gpis_c3(args.ntrain, 2048, threads)
            line["cpu_baseline"] = {"value": v, "unit": "queries/s", "cores": threads, "kind": "port",
                                    "sample": "2048 of the queries: one pred (LU of the 4096^2 matrix + MxM covariance) + "
                                              "one compute_normal (explicit inverse), %.1f s" % dt}
    elif workload == "sdf":
        from bench_sdf import cpu_sdf, run_bench_sdf
        line = run_bench_sdf(args, rank, world, local, ClockSampler, barrier, max_over_ranks, fp64_pipe_peaks, ncu_traffic)
        if rank == 0 and not args.no_cpu_baseline:
            v, dt = cpu_sdf(1000, args.subdiv, 1)
            line["cpu_baseline"] = {"value": v, "unit": "queries/s", "cores": 1, "kind": "port",
                                    "sample": "1000 of the points against the same mesh through the numpy port of the reference "
                                              "kernel (the reference has no CPU path for this op), %.1f s" % dt}
    else:
        from bench_pregrasp import run_bench_c2
        line = run_bench_c2(args, rank, world, local, ClockSampler, barrier, max_over_ranks, fp64_pipe_peaks, measured_peaks, ncu_traffic)
        if rank == 0 and not args.no_cpu_baseline:
            from bench_pregrasp import cpu_pregrasp
            cpu_pregrasp(16, 2, threads)                          # warm-up (torch thread pool, LAPACK)
            v, dt = cpu_pregrasp(64, 10, threads)
            line["cpu_baseline"] = {"value": v, "unit": "candidate-iters/s", "cores": threads, "kind": "port",
                                    "sample": "64 candidates x 10 iterations of the same workload through the oracle port of "
                                              "the reference (torch CPU, per-call LU/inverse GPIS as gpis.py:53,77), %.1f s" % dt}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
