"""Helpers shared by bench.py's workloads: clock sampling inside the timed region, measured peaks, rank plumbing."""
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))


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
    """DRAM bytes (read + write) per launch of a kernel from a committed ncu --set full summary under profiles/
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


_PIPES = {}


def pipe_peaks():
    """FP64 (DFMA / DMMA) and tcgen05 kind::i8 rates are not in MEASURED_PEAKS.json; measured here, once per process, with
    the repo's own microbenchmark (csrc/cdx_pipebench.cu)."""
    if "v" in _PIPES:
        return _PIPES["v"]
    exe = os.path.join(ROOT, "compliancedex_b200", "csrc", "build", "cdx_pipebench")
    try:
        env = dict(os.environ)
        if "LOCAL_RANK" in env and torch.cuda.is_available():
            env["CUDA_VISIBLE_DEVICES"] = str(_physical_index(int(env["LOCAL_RANK"])))
        out = subprocess.run([exe], capture_output=True, text=True, timeout=180, env=env).stdout.strip().splitlines()[-1]
        _PIPES["v"] = json.loads(out)
    except Exception as e:  # noqa: BLE001
        _PIPES["v"] = {"error": str(e)}
    return _PIPES["v"]


def _physical_index(local):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        parts = [v.strip() for v in vis.split(",")]
        if local < len(parts):
            return parts[local]
    return local


def i8_peak(pipes, sustained=False):
    """best measured tcgen05.mma kind::i8 rate (TOP/s) over the operand shapes cdx_pipebench times: the burst figures (a kernel
    timed alone) or the sustained ones (the same loop back to back for ~3 s under the power cap: a kernel timed inside a long step)"""
    vals = [v for k, v in pipes.items() if k.startswith("i8_mma_") and k.endswith("_sustained") == sustained
            and isinstance(v, (int, float)) and v > 0]
    return max(vals) if vals else None


def ranks():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def dist_setup():
    rank, world, local = ranks()
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        # NCCL prints its version banner to STDOUT when the first communicator is created: send fd 1 to stderr while the process
        # group comes up (init + one warm-up collective), so that stdout carries exactly one JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local))
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


def timed(step, steps, world, rank, local, sample_clocks=True):
    """K steps bracketed by barrier + synchronize on both sides, CUDA events on the current stream, max over ranks.
    -> (ms per step, clocks dict or None)"""
    sampler = ClockSampler(local) if (sample_clocks and rank == 0) else None
    barrier(world)
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier(world)
    local_ms = e0.elapsed_time(e1)
    ms = max_over_ranks(local_ms, world)
    if world > 1:
        sys.stderr.write("[bench] rank %d: %.3f ms per step on this rank, %.3f max over ranks\n" % (rank, local_ms / steps, ms / steps))
    return ms / steps, (sampler.stop() if sampler else None)


def reference_kind():
    """"reference": the reference's own Python (in /root/reference, or its unmodified copy oracle/_ref/reference_py on the
    GPU box) runs on the host cores; "port": only the oracle restatement is available."""
    from oracle import ref_harness as rh
    return "reference" if rh.available() else "port"


def cpu_baseline_subprocess(workload, extra=()):
    """The cpu_baseline leg: the reference arm of the same workload in a child process with CUDA hidden (the reference
    harness patches torch for CPU execution - kept out of this process).  One bounded step."""
    warm = "1" if workload.startswith("pregrasp") or workload == "gpis_c3" else "0"    # cheap steps: keep import / first-call costs out
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", workload, "--steps", "1",
           "--warmup", warm, "--secondary", "none"] + list(extra)
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    env["CUDA_VISIBLE_DEVICES"] = ""
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
        line = json.loads(out.stdout.strip().splitlines()[-1])
        return line["cpu_baseline"]
    except Exception as e:  # noqa: BLE001
        return {"value": None, "unit": None, "cores": os.cpu_count(), "kind": "port", "sample": "cpu baseline failed: %r" % (e,)}
