#!/usr/bin/env python
"""Gradient all-reduce bandwidth sweep, 4 KB - 256 MB (BASELINE.json config 5).

Launch with torchrun (one rank per GPU).  For every size it times, on the device (CUDA events, max
over ranks):
  fused-sgd   b200cnn_comm_allreduce_sgd   fused all-reduce + 1/W + SGD (one-shot / two-shot / NVLS)
  fused-ar    b200cnn_comm_allreduce       the same kernels without the update
  nccl        dist.all_reduce              library all-reduce only
  nccl+sgd    dist.all_reduce + scale/SGD elementwise kernel (the designated baseline path)
and prints one JSON line per (size, path).  Bandwidths: algbw = bytes / t, busbw = algbw * 2(W-1)/W (the NCCL
convention, for comparison with nccl-tests only).  Roofline: `link_bytes` = bytes that must cross THIS GPU's NVLink
ports in the busier direction for that path, `frac_of_link` = link_bytes / 770 GB/s (measured peer copy,
B200_PROFILING.md) / t.  Per path (N = message bytes, W ranks):
  one_shot      (W-1) N          every rank reads every peer's whole slice
  two_shot      2 (W-1)/W N      reads its shard from W-1 peers, receives W-1 shards of the result
  nvls          N                in: N/W reduced by the switch + (W-1)/W N of results; out: (W-1)/W N + N/W
  ll_one_shot   2 (W-1) N        packets are half payload, half tag
  ll_two_shot   3 (W-1)/W N      tagged shards out, plain results out
  nccl          N                assumes its best case on this box (NVLS); fractions above are therefore never > 1
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

from mpi_cuda_cnn_b200.ops import functional as F200
from mpi_cuda_cnn_b200.parallel import FusedComm, init_distributed

LINK_GBS = 770.0


def timed(fn, iters, stream):
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    start.record(stream)
    for _ in range(iters):
        fn()
    end.record(stream)
    stream.synchronize()
    t = torch.tensor([start.elapsed_time(end) / iters], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e-3


def link_bytes(path, nbytes, world, multicast):
    """Bytes through this GPU's NVLink ports in the busier direction (module docstring)."""
    algo = path.split("/")[1] if "/" in path else "nccl"
    w = world
    if algo == "auto":  # what AUTO resolves to (include/b200cnn_comm.h)
        if nbytes * (w - 1) <= (448 << 10):
            algo = "ll_one_shot"
        elif multicast and w >= 4 and nbytes >= (256 << 10):
            algo = "nvls"
        elif nbytes // w + 32 <= (1 << 20) and nbytes <= (8 << 20):
            algo = "ll_two_shot"
        else:
            algo = "nvls" if (multicast and nbytes >= (8 << 20)) else "two_shot"
    return {"one_shot": (w - 1) * nbytes, "two_shot": 2 * (w - 1) / w * nbytes, "nvls": nbytes,
            "ll_one_shot": 2 * (w - 1) * nbytes, "ll_two_shot": 3 * (w - 1) / w * nbytes, "nccl": nbytes}[algo]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-bytes", type=int, default=4096)
    ap.add_argument("--max-bytes", type=int, default=256 << 20)
    ap.add_argument("--algos", default="auto,one_shot,two_shot,nvls,ll_one_shot,ll_two_shot")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    rank, world, device = init_distributed()
    max_floats = args.max_bytes // 4
    comm = FusedComm(max_floats, rank, world, device)
    stream = torch.cuda.Stream()
    results = []
    sizes = []
    b = args.min_bytes
    while b <= args.max_bytes:
        sizes.append(b)
        b *= 4
    # the model's own message: 360,810 fp32 parameters (padded flat buffer of the engine)
    sizes.append(1443840)
    algos = [a for a in args.algos.split(",") if a != "nvls" or comm.has_multicast]
    g = comm.grads(0)
    p = comm.params()
    g.normal_()
    p.zero_()
    nccl_buf = torch.randn(max_floats, device=device)
    nccl_w = torch.zeros(max_floats, device=device)
    with torch.cuda.stream(stream):
        for nbytes in sorted(set(sizes)):
            n = nbytes // 4 // 4 * 4
            iters = 200 if nbytes <= (1 << 20) else (50 if nbytes <= (32 << 20) else 15)
            rows = []
            for algo in algos:
                if algo == "one_shot" and nbytes > (8 << 20):
                    continue
                if algo == "ll_one_shot" and nbytes > (1 << 20):
                    continue
                if algo == "ll_two_shot" and nbytes // world + 32 > (1 << 20):
                    continue
                for name, fn in (("fused-sgd", lambda: comm.allreduce_sgd(1e-6, 0, n, algo=algo, stream=stream)),
                                 ("fused-ar", lambda: comm.allreduce(0, n, algo=algo, stream=stream))):
                    for _ in range(5):
                        fn()
                    t = timed(fn, iters, stream)
                    rows.append((f"{name}/{algo}", t))
            view, wview = nccl_buf[:n], nccl_w[:n]
            for name, fn in (("nccl", lambda: dist.all_reduce(view)),
                             ("nccl+sgd", lambda: (dist.all_reduce(view), F200.sgd_update_(wview, view, 1e-6 / world, zero_grad=False)))):
                for _ in range(5):
                    fn()
                rows.append((name, timed(fn, iters, stream)))
            for name, t in rows:
                alg = n * 4 / t / 1e9
                bus = alg * 2 * (world - 1) / world
                link = link_bytes(name, n * 4, world, comm.has_multicast)
                rec = {"bytes": n * 4, "path": name, "world": world, "us": t * 1e6, "algbw_gbs": alg, "busbw_gbs": bus,
                       "link_bytes": link, "frac_of_link": link / (LINK_GBS * 1e9) / t, "multicast": comm.has_multicast,
                       "region": comm.region.kind}
                results.append(rec)
                if rank == 0:
                    print(json.dumps(rec), flush=True)
    if comm.error():
        raise SystemExit("communicator reported a peer timeout")
    if rank == 0 and args.out:
        with open(args.out, "w") as f:
            json.dump(results, f, indent=1)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
