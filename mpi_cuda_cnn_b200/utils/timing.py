"""Device-side timing helpers and the clock / throttle-reason sampler used by the benchmarks."""
import statistics
import subprocess
import threading
import time

import torch

_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Polls nvidia-smi in a background thread while a timed region runs."""

    def __init__(self, gpu_index=0, period_s=0.1):
        self.gpu_index = gpu_index
        self.period_s = period_s
        self.samples = []
        self._stop = threading.Event()
        self._thread = None

    def _poll(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits"],
                                     stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([f.strip() for f in out.splitlines()[0].split(",")])
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
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx.append(float(s[1]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, s[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


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
