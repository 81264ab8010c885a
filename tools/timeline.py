"""Kernel timeline of ONE steady-state training step (diagnosis only; numbers under a profiler are not bench values).

Runs the graph-captured step under torch.profiler (Kineto/CUPTI records the kernels our .so launches too), takes
the kernels of one step in the middle of the run and prints start offset, duration, stream and the gap to the
previous kernel on the same stream.  Usage: python tools/timeline.py [batch] [steps] [--pool] [--simt] [--eager]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

from mpi_cuda_cnn_b200.engine import Engine
from mpi_cuda_cnn_b200.ops import functional as F200

POOL, SIMT, EAGER = "--pool" in sys.argv, "--simt" in sys.argv, "--eager" in sys.argv
_pos = [a for a in sys.argv[1:] if not a.startswith("--")]
B = int(_pos[0]) if len(_pos) > 0 else 256
STEPS = int(_pos[1]) if len(_pos) > 1 else 40
if SIMT:
    F200.set_backend("simt")
WORLD = int(os.environ.get("WORLD_SIZE", "1"))
RANK = int(os.environ.get("RANK", "0"))
dev = torch.device("cuda:0")
if WORLD > 1:  # under torchrun: every rank trains, every rank prints its own view (timeline_dp: rank 0 and the last rank matter)
    from mpi_cuda_cnn_b200.parallel import DataParallelTrainer, init_distributed
    RANK, WORLD, dev = init_distributed()
    _tr = DataParallelTrainer(RANK, WORLD, dev, batch=B, mode="fused")
    eng = _tr.engine
    if RANK not in (0, WORLD - 1):
        sys.stdout = open(os.devnull, "w")
    else:
        sys.stdout = open(os.environ.get("TIMELINE_OUT", "timeline") + f".rank{RANK}.txt", "w")
elif os.environ.get("B200CNN_BENCH_FORCE_DP") == "1":  # the data-parallel launch structure with a world of one
    from mpi_cuda_cnn_b200.parallel import DataParallelTrainer
    _tr = DataParallelTrainer(0, 1, dev, batch=B, mode="fused")
    eng = _tr.engine
else:
    from mpi_cuda_cnn_b200 import mnist_layers, pool_layers
    eng = Engine(pool_layers() if POOL else mnist_layers(), max_batch=B, device=dev)
    eng.init_params(seed=1)
x, y = F200.synthetic_batch(B * 8, seed=1, device=dev)
xs = [x[i * B:(i + 1) * B].contiguous() for i in range(8)]
ys = [y[i * B:(i + 1) * B].contiguous() for i in range(8)]
stream = torch.cuda.Stream(device=dev)
with torch.cuda.stream(stream):
    for i in range(20):
        eng.train_step(xs[i % 8], ys[i % 8], 0.1 / B, stream=stream, use_graph=not EAGER)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(STEPS):
            eng.train_step(xs[i % 8], ys[i % 8], 0.1 / B, stream=stream, use_graph=not EAGER)
        torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "emcpy" not in e.name and "emset" not in e.name]
ev.sort(key=lambda e: e.time_range.start)
per = len(ev) // STEPS
print(f"{len(ev)} kernels in {STEPS} steps -> {per} per step")
k0 = per * (STEPS // 2)
step = ev[k0:k0 + per]
t0 = step[0].time_range.start
last_end = {}
print(f"{'start':>8} {'dur':>7} {'gap':>7}  stream  kernel")
for e in step:
    s, d = e.time_range.start - t0, e.time_range.end - e.time_range.start
    sid = getattr(e, "stream", None)
    if sid is None:
        sid = getattr(e, "device_resource_id", -1)
    gap = s - last_end[sid] if sid in last_end else float("nan")
    last_end[sid] = s + d
    print(f"{s:8.2f} {d:7.2f} {gap:7.2f}  {sid!s:>6}  {e.name[:90]}")
nxt = ev[k0 + per].time_range.start - t0 if k0 + per < len(ev) else float("nan")
print(f"next step starts at {nxt:.2f} us; busy (sum of durations) {sum(e.time_range.end - e.time_range.start for e in step):.2f} us")
sys.stdout.flush()
if WORLD > 1:
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()
