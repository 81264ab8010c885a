#!/usr/bin/env python
"""Headline benchmark: training images/s of the MNIST-shaped CNN at per-GPU batch 256 (BASELINE.json).

    python bench.py [--gpus N --steps K --warmup W]            this framework (N>1: launch with torchrun)
    python bench.py --impl nccl ...                            same engine, NCCL all-reduce + scale/SGD kernel update
    python bench.py --impl reference [--gpus N ...]            the unmodified reference from baseline/_ref

One JSON line on stdout (rank 0).  `value` is the device-timed whole-job throughput: CUDA events around exactly K
steps, host barrier + synchronize on both sides, a DEVICE-side rendezvous of all ranks on the timed stream right
before the first event (so no rank's clock starts while another rank is still getting there), max over ranks.
`e2e` is the same metric through the public API with the host->device copy of every step's inputs (pinned memory)
and a device->host read of every step's loss inside the timed region.  `baselines` carries the same-precision bar
measured in the same process at the same N/K/W (NCCL all-reduce + separate scale/SGD kernel over the same
forward/backward), `parity` the check of the GPU engine's training metric against the fp64 CPU engine (the chain
back to /root/reference/cnn.c, tests/test_cpu_engine.py) — the process exits non-zero when that check fails.
"""
import argparse
import json
import os
import struct
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 256  # per GPU (BASELINE.json configs 3 and 4)
RATE = 0.1   # cnn.c:446
MODEL = "mnist-cnn 1x28x28 -> conv3x3s2(16)+relu -> conv3x3s2(32)+relu -> fc200 tanh -> fc200 tanh -> fc10 softmax (360,810 params)"
MODEL_POOL = "mnist-cnn 1x28x28 -> conv3x3(16)+relu -> maxpool2 -> conv3x3(32)+relu -> maxpool2 -> fc200 tanh -> fc200 tanh -> fc10 softmax"
# Serial cnn.c measured by the survey on the authoring machine: the only published-equivalent number.
BASELINE_IMG_S = 113.0
PARITY_TOL = 5e-3  # |GPU TF32 - CPU fp64| of the reference's per-step 'error' metric (mean squared p - y)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "nccl"])
    ap.add_argument("--algo", default="auto", choices=["auto", "one_shot", "two_shot", "nvls", "ll_one_shot", "ll_two_shot"])
    ap.add_argument("--backend", default="auto", choices=["auto", "simt", "tcgen05"])
    ap.add_argument("--pool", action="store_true", help="the conv -> relu -> max-pool variant of the topology")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-baseline", action="store_true", help="skip the in-process NCCL/unfused baseline arm")
    ap.add_argument("--parity-steps", type=int, default=50, help="mini-batches of the GPU-vs-fp64-CPU check (0: skip)")
    ap.add_argument("--dataset", type=int, default=204800, help="device-resident synthetic samples per rank (x784 B; default > L2)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# reference arm — nothing of this repository's package is imported on this path
# ------------------------------------------------------------------------------------------------
def _write_idx(path, dims, payload):
    """IDX file as /root/reference/cnn.c:352-383 parses it: 00 00 08 <ndims>, big-endian dims, uint8 payload."""
    with open(path, "wb") as f:
        f.write(bytes([0, 0, 8, len(dims)]))
        for d in dims:
            f.write(struct.pack(">I", d))
        f.write(payload)


def _synthetic_idx(directory, ntrain, ntest):
    """MNIST-shaped learnable data in plain Python/numpy: one random prototype per class plus pixel noise."""
    import numpy as np

    os.makedirs(directory, exist_ok=True)
    rng = np.random.default_rng(1)
    protos = rng.integers(0, 256, size=(10, 28, 28))
    paths = []
    for name, n in (("train", ntrain), ("t10k", ntest)):
        y = rng.integers(0, 10, size=n)
        x = np.clip(protos[y] // 2 + rng.integers(0, 128, size=(n, 28, 28)), 0, 255).astype(np.uint8)
        paths += [os.path.join(directory, f"{name}-images-idx3-ubyte"), os.path.join(directory, f"{name}-labels-idx1-ubyte")]
        _write_idx(paths[-2], (n, 28, 28), x.tobytes())
        _write_idx(paths[-1], (n,), y.astype(np.uint8).tobytes())
    return paths


def run_reference(args):
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    ref = os.path.join(ROOT, "baseline", "_ref")
    binary = os.path.join(ref, "cnn" if world == 1 else "cnnmpi")
    if not os.path.exists(binary) and os.path.isdir(os.path.join(ref, "src")):
        subprocess.run(["bash", os.path.join(ROOT, "baseline", "install_reference.sh"), os.path.join(ref, "src")],
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if not os.path.exists(binary):
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref has no built reference (run baseline/install_reference.sh); "
                              "the reference's CUDA programs do not compile at all"}))
        return 0

    # The reference has no flags: 10 epochs, batch 32, fp64, one sample at a time (cnn.c:446-451).
    # K steps of OUR batch would be K*256 samples; the serial reference needs ~9 ms per sample, so it gets
    # a fixed, smaller budget: train_size samples x 10 epochs.
    train_size = 160 * max(world, 1)
    tmp = os.environ.get("B200CNN_REF_TMP") or os.path.join(tempfile.gettempdir(), f"b200cnn_ref_{os.environ.get('MASTER_PORT', '0')}")
    if rank == 0:
        paths = _synthetic_idx(tmp, train_size, 10)
        open(os.path.join(tmp, "ready"), "w").close()
    else:
        while not os.path.exists(os.path.join(tmp, "ready")):
            time.sleep(0.05)
        paths = [os.path.join(tmp, n) for n in ("train-images-idx3-ubyte", "train-labels-idx1-ubyte", "t10k-images-idx3-ubyte", "t10k-labels-idx1-ubyte")]
    t0 = time.perf_counter()
    res = subprocess.run([binary] + paths, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
    dt = time.perf_counter() - t0
    if rank == 0:
        samples = 10 * (train_size // world) * world
        img_s = samples / dt
        out = {
            "impl": "reference", "metric": "train_images_per_sec", "value": img_s, "unit": "images/s", "n_gpus": args.gpus,
            "steps": samples // 32, "warmup": 0, "ms_per_step": 1e3 * dt / max(samples // 32, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": img_s / BASELINE_IMG_S, "dtype": "fp64", "data": "synthetic",
            "exit_code": res.returncode,
            "config": {"model": MODEL, "global_batch": 32 if world == 1 else world, "seq_len": None,
                       "parallelism": "none (serial cnn.c)" if world == 1 else f"mpi-dp{world} (cnnmpi.c over an in-repo shared-memory mpi.h shim; no MPI in the image)",
                       "program": os.path.basename(binary), "build": "gcc, reference Makefile flags (no -O)",
                       "samples": samples, "wall_s": dt},
            # A CPU program fed from files on the host: its wall-clock rate IS its end-to-end rate (data load, training
            # and the 10-image test pass included; nothing crosses PCIe).
            "e2e": {"value": img_s, "unit": "images/s", "ms_per_step": 1e3 * dt / max(samples // 32, 1),
                    "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "The reference's GPU programs (CUDAcnn.cu, CUDAMPI.c/.cu) do not compile (baseline/_ref/BUILD_STATUS.txt); "
                    "this is its unmodified CPU trainer, which uses no GPU. Wall-clock timed (the reference has no timers); "
                    "cnnmpi.c never reads the IDX payload and all-reduces per sample (SURVEY D6-D10), so only its speed is meaningful. "
                    "fp64 on the CPU at batch 32 against TF32 on the GPU at batch 256 is not a like-for-like ratio; the "
                    "same-precision bar is the `baselines` block of the b200 arm.",
        }
        print(json.dumps(out))
    return 0


# ------------------------------------------------------------------------------------------------
# this framework
# ------------------------------------------------------------------------------------------------
class _Arm:
    """One update path (fused NVLink all-reduce+SGD, or NCCL + elementwise) over its own engine replica."""

    def __init__(self, mode, args, rank, world, device, layers):
        import torch

        from mpi_cuda_cnn_b200 import Engine
        from mpi_cuda_cnn_b200.parallel import DataParallelTrainer

        self.mode, self.world, self.device = mode, world, device
        self.use_graph = not args.no_graph
        if world == 1 and mode == "fused" and os.environ.get("B200CNN_BENCH_FORCE_DP") != "1":
            self.trainer = None
            self.engine = Engine(layers, max_batch=BATCH, device=device)
            self.engine.init_params(seed=0)
        else:
            self.trainer = DataParallelTrainer(rank, world, device, layers=layers, batch=BATCH, mode=mode, algo=args.algo)
            self.engine = self.trainer.engine
        self._tiny = torch.zeros(1, device=device)

    @property
    def comm(self):
        return self.trainer.comm if self.trainer is not None else None

    def step(self, x, y):
        if self.trainer is None:
            self.engine.train_step(x, y, RATE / BATCH, use_graph=self.use_graph)
        else:
            self.trainer.train_step(x, y, RATE, use_graph=self.use_graph)

    def rendezvous(self, stream):
        """Device-side meeting point of all ranks, enqueued on `stream`: whatever follows on the stream starts on
        every GPU only after the slowest rank's stream has got here."""
        import torch.distributed as dist

        if self.world == 1:
            return
        if self.comm is not None:
            self.comm.barrier(stream)  # one flag exchange over NVLink inside a kernel on `stream`
        else:
            dist.all_reduce(self._tiny)  # NCCL kernel ordered into the current stream

    def close(self):
        if self.trainer is not None:
            self.trainer.close()
        else:
            self.engine.close()


def _host_barrier(world):
    import torch
    import torch.distributed as dist

    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def _device_timed(arm, F200, stream, data, rank, K, W, clocks=None):
    """W untimed + K timed steps; every step trains on a random batch gathered on a second stream from a device-resident
    set larger than L2 (cnn.c:455: rand() % train_size).  Returns elapsed ms of the K steps on this rank."""
    import torch

    device = arm.device
    data_x, data_y = data
    stage = [(torch.empty((BATCH, 784), dtype=torch.uint8, device=device), torch.empty((BATCH,), dtype=torch.uint8, device=device))
             for _ in range(2)]
    gather_stream = torch.cuda.Stream(device=device)
    gathered = [torch.cuda.Event(), torch.cuda.Event()]
    freed = [torch.cuda.Event(), torch.cuda.Event()]
    for ev in freed:
        ev.record(stream)

    def gather(i):  # batch i+1 is gathered while step i runs (one captured graph per staging buffer)
        with torch.cuda.stream(gather_stream):
            gather_stream.wait_event(freed[i % 2])
            F200.gather_batch(data_x, data_y, BATCH, seed=3, step=i, out=stage[i % 2])
            gathered[i % 2].record(gather_stream)

    def step(i):
        stream.wait_event(gathered[i % 2])
        arm.step(*stage[i % 2])
        freed[i % 2].record(stream)

    gather(0)
    for i in range(W):
        gather(i + 1)
        step(i)
    arm.engine.reset_metrics()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _host_barrier(arm.world)
    # Nothing rank-variable on the host from here to ev0; the ranks' streams meet on the device first.
    arm.rendezvous(stream)
    ev0.record(stream)
    if clocks is not None:
        clocks.begin()
    for i in range(K):
        if i + 1 < K:
            gather(W + i + 1)
        step(W + i)
    ev1.record(stream)
    if clocks is not None:
        clocks.sample_now()  # the GPU is still working through the queued steps: a sample under load
    stream.synchronize()
    if clocks is not None:
        clocks.end()
    _host_barrier(arm.world)
    gather_stream.synchronize()
    return ev0.elapsed_time(ev1)


def _e2e_timed(arm, stream, host_x, host_y, K, W):
    """K steps through the public host-fed API: pinned host inputs -> device every step, loss read back every step."""
    import torch

    device, engine, mode = arm.device, arm.engine, arm.mode
    pool = host_x.shape[0]
    # Two pinned result slots: the host reads step i-1's loss while step i runs (each step's inputs are still
    # copied from pinned host memory and each step's loss is still read back inside the timed region).
    loss_host = [torch.zeros(4, dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ready = [torch.cuda.Event(), torch.cuda.Event()]
    metrics = engine.metrics_tensor()
    stage = [(torch.empty((BATCH, 784), dtype=torch.uint8, device=device), torch.empty((BATCH,), dtype=torch.uint8, device=device))
             for _ in range(2)]
    copy_stream = torch.cuda.Stream(device=device)
    copied = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    for ev in consumed:
        ev.record(stream)

    def upload(i):
        if mode == "fused":
            return  # train_step_host copies on the engine's own copy stream
        sx, sy = stage[i % 2]
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[i % 2])  # the step that last read this buffer has finished
            sx.copy_(host_x[i % pool], non_blocking=True)
            sy.copy_(host_y[i % pool], non_blocking=True)
            copied[i % 2].record(copy_stream)

    def enqueue(i):
        if mode == "fused":
            # the public host-fed step: H2D of this step's inputs (engine's copy stream, double-buffered), the
            # captured step, D2H of the loss — one native call, nothing blocks
            engine.train_step_host(host_x[i % pool], host_y[i % pool], RATE / BATCH, metrics_out=loss_host[i % 2], stream=stream)
        else:
            sx, sy = stage[i % 2]
            stream.wait_event(copied[i % 2])
            arm.step(sx, sy)
            consumed[i % 2].record(stream)
            loss_host[i % 2].copy_(metrics, non_blocking=True)
        loss_ready[i % 2].record(stream)

    def collect(i):
        loss_ready[i % 2].synchronize()
        return float(loss_host[i % 2][0])

    for i in range(W):
        upload(i)
        enqueue(i)
        collect(i)
    engine.reset_metrics()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    losses = []
    _host_barrier(arm.world)
    arm.rendezvous(stream)
    e0.record(stream)
    upload(0)
    for i in range(K):
        if i + 1 < K:
            upload(i + 1)
        enqueue(i)
        if i > 0:
            losses.append(collect(i - 1))
    losses.append(collect(K - 1))
    e1.record(stream)
    stream.synchronize()
    _host_barrier(arm.world)
    return e0.elapsed_time(e1), losses


def _parity_check(layers, device, steps):
    """The GPU engine (TF32) and the fp64 CPU engine train on IDENTICAL mini-batches from the same initial weights;
    compares the reference's training metric (mean over the batch of mean_j (p_j - y_j)^2, cnn.c:275-282) step by step."""
    import numpy as np
    import torch

    from mpi_cuda_cnn_b200 import Engine
    from mpi_cuda_cnn_b200.utils import make_synthetic

    x, y = make_synthetic(BATCH * steps, seed=7)
    eng = Engine(layers, max_batch=BATCH, device=device)
    cpu = eng.init_params(seed=0)
    dx, dy = torch.from_numpy(x.reshape(-1, 784)).to(device), torch.from_numpy(y).to(device)
    gpu_curve, cpu_curve = [], []
    xs, ys = torch.empty((BATCH, 784), dtype=torch.uint8, device=device), torch.empty((BATCH,), dtype=torch.uint8, device=device)
    stream = torch.cuda.Stream(device=device)
    with torch.cuda.stream(stream):
        for s in range(steps):
            xs.copy_(dx[s * BATCH:(s + 1) * BATCH])
            ys.copy_(dy[s * BATCH:(s + 1) * BATCH])
            eng.reset_metrics()
            eng.train_step(xs, ys, RATE / BATCH, use_graph=True)
            gpu_curve.append(eng.read_metrics()["mse_sum"] / BATCH)
    eye = np.eye(10)
    batched = cpu.batch_supported
    for s in range(steps):
        xs = x[s * BATCH:(s + 1) * BATCH].reshape(BATCH, -1).astype(np.float64) / 255.0
        ts = eye[y[s * BATCH:(s + 1) * BATCH]]
        if batched:
            err = cpu.learn_batch(xs, ts)
        else:
            err = np.empty(BATCH)
            for i in range(BATCH):
                cpu.set_inputs(xs[i])
                cpu.learn_outputs(ts[i])
                err[i] = cpu.error_total()
        cpu.update(RATE / BATCH)
        cpu_curve.append(float(err.mean()))
    g, c = np.array(gpu_curve), np.array(cpu_curve)
    pg, pc = eng.get_params().astype(np.float64), cpu.get_params()
    eng.close()
    diff = float(np.abs(g - c).max())
    return {"metric_max_abs_diff": diff, "tol": PARITY_TOL, "ok": bool(diff <= PARITY_TOL), "steps": steps, "batch": BATCH,
            "gpu_last": float(g[-1]), "cpu_fp64_last": float(c[-1]), "gpu_first": float(g[0]), "cpu_fp64_first": float(c[0]),
            "param_max_rel_diff": float(np.abs(pg - pc).max() / np.abs(pc).max()),
            "oracle": "native fp64 CPU engine (bit-identical to /root/reference/cnn.c in its quirks mode)"}


def run_b200(args):
    import torch
    import torch.distributed as dist

    from mpi_cuda_cnn_b200 import mnist_layers, pool_layers
    from mpi_cuda_cnn_b200.ops import functional as F200
    from mpi_cuda_cnn_b200.parallel import init_distributed
    from mpi_cuda_cnn_b200.utils import ClockSampler

    rank, world, device = init_distributed()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")
    F200.set_backend(args.backend)
    layers = pool_layers() if args.pool else mnist_layers()
    mode = "nccl" if args.impl == "nccl" else "fused"
    K, W = args.steps, max(args.warmup, 3)
    n_data = max(args.dataset, BATCH)
    # NVML init + handle lookup happen here, long before any timed region (they take milliseconds and differ per rank).
    clocks = ClockSampler(gpu_index=device.index or 0, period_s=0.001)

    stream = torch.cuda.Stream(device=device)
    arm = _Arm(mode, args, rank, world, device, layers)
    engine = arm.engine
    with torch.cuda.stream(stream):
        # Device-resident synthetic shard for this rank (K10), larger than L2 so successive steps read different lines.
        data = F200.synthetic_batch(n_data, seed=1, first_index=rank * n_data, device=device)
        with clocks:
            ms = _device_timed(arm, F200, stream, data, rank, K, W, clocks)
        m_dev = engine.read_metrics()

        pool = max(1, min(64, n_data // BATCH))
        host_x = torch.empty((pool, BATCH, 784), dtype=torch.uint8).pin_memory()
        host_y = torch.empty((pool, BATCH), dtype=torch.uint8).pin_memory()
        host_x.copy_(data[0][: pool * BATCH].reshape(pool, BATCH, 784).cpu())
        host_y.copy_(data[1][: pool * BATCH].reshape(pool, BATCH).cpu())
        ms_e2e, losses = _e2e_timed(arm, stream, host_x, host_y, K, W)
        comm_error = bool(arm.comm is not None and arm.comm.error())
        launches_per_step = engine.launches_per_step + 1  # + the batch-gather kernel
        region_kind = arm.comm.region.kind if arm.comm is not None else None
        has_mc = arm.comm.has_multicast if arm.comm is not None else None
        arm.close()

        # Same-precision bar, same process, same N / K / W: identical graph-captured forward+backward, update by
        # NCCL all-reduce + one scale/SGD kernel (at N=1: the un-fused update, no collective).
        ms_base = None
        if mode == "fused" and not args.no_baseline:
            base = _Arm("nccl", args, rank, world, device, layers)
            ms_base = _device_timed(base, F200, stream, data, rank, K, W)
            base.close()

    vals = [ms, ms_e2e, ms_base if ms_base is not None else 0.0]
    t = torch.tensor(vals, dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e, ms_base_max = t.tolist()
    if comm_error:
        raise SystemExit("fused communicator reported a peer timeout")
    parity = None
    if rank == 0 and args.parity_steps > 0:
        parity = _parity_check(layers, device, args.parity_steps)
    value = world * BATCH * K / (ms / 1e3)
    e2e = world * BATCH * K / (ms_e2e / 1e3)
    rc = 0
    if rank == 0:
        ck = clocks.summary()
        tc = F200_backend_uses_tc(args.backend)
        out = {
            "metric": "train_images_per_sec", "value": value, "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": value / BASELINE_IMG_S,
            "dtype": "tf32" if tc else "fp32",
            "dtype_note": ("TF32 multiply / fp32 accumulate on tcgen05, fp32 master weights; the reference is fp64 — see `parity` "
                           f"(tolerance {PARITY_TOL} on the reference's training metric)") if tc else "fp32 CUDA cores",
            "data": "synthetic", "impl": args.impl,
            "config": {"model": MODEL_POOL if args.pool else MODEL, "global_batch": BATCH * world, "seq_len": None, "parallelism": f"dp{world}",
                       "per_gpu_batch": BATCH, "update": ("fused p2p allreduce+sgd (%s)" % args.algo) if mode == "fused" else "nccl allreduce + scale/sgd kernel",
                       "cuda_graph": not args.no_graph, "backend": args.backend, "region": region_kind, "nvls_multicast": has_mc,
                       "l2_policy": f"inputs larger than L2: each step gathers a random batch from a {n_data * 784 / 2**20:.0f} MiB device-resident set",
                       "timing": "cuda events on the launching stream; host barrier, then device-side rendezvous of all ranks, then ev0",
                       "final_mse_metric": m_dev["mse_sum"] / max(m_dev["count"], 1), "final_accuracy": m_dev["correct"] / max(m_dev["count"], 1)},
            "clocks": {"sm_mhz": ck["sm_mhz"], "sm_max_mhz": ck["sm_max_mhz"], "reasons": ck["reasons"], "samples": ck["samples"]},
            "e2e": {"value": e2e, "unit": "images/s", "ms_per_step": ms_e2e / K, "h2d_bytes_per_step": BATCH * 784 + BATCH,
                    "d2h_bytes_per_step": 16,
                    "last_step_mean_ce": (losses[-1] - losses[-2]) / BATCH if len(losses) > 1 else None},
            "gpu_launches": launches_per_step * K if mode == "fused" else None,
            "launches_per_step": launches_per_step,
        }
        if ms_base is not None:
            out["baselines"] = {"what": "same graph-captured forward+backward; update = NCCL all-reduce + one scale/SGD kernel" if world > 1
                                else "same graph-captured forward+backward; update = one separate SGD kernel (no collective at N=1)",
                                "nccl_ms_per_step": ms_base_max / K, "nccl_images_per_sec": world * BATCH * K / (ms_base_max / 1e3),
                                "fused_over_nccl": ms_base_max / ms, "dtype": out["dtype"], "steps": K, "warmup": W}
        if parity is not None:
            out["parity"] = parity
            if not parity["ok"]:
                rc = 3
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return rc


def F200_backend_uses_tc(backend):
    from mpi_cuda_cnn_b200 import _native as N

    info = N.cuda().b200cnn_build_info().decode()
    return backend != "simt" and "tcgen05-gemm" in info


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
