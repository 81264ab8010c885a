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

    NVML (a ~1 ms call) when available, else `nvidia-smi` (tens of ms per sample).

    Construct it (NVML init, handle lookup: milliseconds, and different on every rank) and enter it BEFORE the barrier
    that precedes a timed region, never between the barrier and the first event.  `begin()` / `end()` mark the timed
    window with host timestamps; `summary()` then reports only the samples taken inside it.  `sample_now()` takes one
    sample on the calling thread (NVML releases the GIL; call it right after the timed launches have been enqueued,
    while the GPU is still working through them) so that even a 1 ms region has a sample taken under load."""

    _REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown"}

    def __init__(self, gpu_index=0, period_s=0.002):
        self.gpu_index = gpu_index
        self.period_s = period_s
        self.samples = []  # (sm_mhz, sm_max_mhz, set(reasons), host time)
        self._t0 = self._t1 = None
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
        self.samples.append((sm, self._max, {name for bit, name in self._REASONS.items() if mask & bit}, time.perf_counter()))

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits"],
                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=5).stdout.strip()
        if not out:
            return
        f = [v.strip() for v in out.splitlines()[0].split(",")]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        self.samples.append((float(f[0]), float(f[1]), {n for n, v in zip(names, f[3:7]) if v.lower().startswith("active")},
                             time.perf_counter()))

    def sample_now(self):
        try:
            self._sample_nvml() if self._nvml else self._sample_smi()
        except Exception:
            pass

    def _poll(self):
        while not self._stop.is_set():
            self.sample_now()
            self._stop.wait(self.period_s)

    def begin(self):
        self._t0 = time.perf_counter()

    def end(self):
        self._t1 = time.perf_counter()

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
        inside = self.samples
        if self._t0 is not None:
            t1 = self._t1 if self._t1 is not None else float("inf")
            inside = [s for s in self.samples if self._t0 <= s[3] <= t1] or self.samples[-1:]
        sm = [s[0] for s in inside]
        mx = [s[1] for s in inside]
        reasons = set().union(*[s[2] for s in inside]) if inside else set()
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
