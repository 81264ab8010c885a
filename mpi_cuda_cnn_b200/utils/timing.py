"""Device-side timing helpers and the clock / throttle-reason sampler used by the benchmarks."""
import statistics
import subprocess
import threading
import time

import torch

_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples SM clock and throttle reasons in a background thread while a timed region runs.

    NVML (a ~1 ms call) when available, else `nvidia-smi` (tens of ms per sample)."""

    _REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown"}

    def __init__(self, gpu_index=0, period_s=0.002):
        self.gpu_index = gpu_index
        self.period_s = period_s
        self.samples = []  # (sm_mhz, sm_max_mhz, set(reasons))
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nvml = pynvml
            # NVML indexes physical GPUs; honour CUDA_VISIBLE_DEVICES when it is a plain index list.
            import os

            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = gpu_index
            if vis and all(v.strip().isdigit() for v in vis.split(",")):
                ids = [int(v) for v in vis.split(",")]
                if gpu_index < len(ids):
                    phys = ids[gpu_index]
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self._max = float(pynvml.nvmlDeviceGetMaxClockInfo(self._handle, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nvml = None

    def _sample_nvml(self):
        n = self._nvml
        sm = float(n.nvmlDeviceGetClockInfo(self._handle, n.NVML_CLOCK_SM))
        try:
            mask = int(n.nvmlDeviceGetCurrentClocksEventReasons(self._handle))
        except Exception:
            mask = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._handle))
        self.samples.append((sm, self._max, {name for bit, name in self._REASONS.items() if mask & bit}))

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits"],
                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=5).stdout.strip()
        if not out:
            return
        f = [v.strip() for v in out.splitlines()[0].split(",")]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        self.samples.append((float(f[0]), float(f[1]), {n for n, v in zip(names, f[3:7]) if v.lower().startswith("active")}))

    def _poll(self):
        while not self._stop.is_set():
            try:
                self._sample_nvml() if self._nvml else self._sample_smi()
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def __enter__(self):
        self._thread = threading.Thread(target=self._poll, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=10)
        return False

    def summary(self):
        sm = [s[0] for s in self.samples]
        mx = [s[1] for s in self.samples]
        reasons = set().union(*[s[2] for s in self.samples]) if self.samples else set()
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvml" if self._nvml else "nvidia-smi"}


def cuda_time_ms(fn, iters, stream=None):
    """Device time of ``iters`` calls of fn() measured with CUDA events on the launching stream."""
    stream = stream or torch.cuda.current_stream()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    start.record(stream)
    for _ in range(iters):
        fn()
    end.record(stream)
    torch.cuda.synchronize()
    return start.elapsed_time(end)


_flush_buf = None


def flush_l2(nbytes=256 << 20):
    """Evict L2 (126 MB on B200) by writing a larger buffer."""
    global _flush_buf
    if _flush_buf is None or _flush_buf.numel() < nbytes:
        _flush_buf = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    _flush_buf.zero_()
