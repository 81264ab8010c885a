#!/usr/bin/env python
"""Headline benchmark: training images/s of the MNIST-shaped CNN at per-GPU batch 256 (BASELINE.json).

    python bench.py [--gpus N --steps K --warmup W]            this framework (N>1: launch with torchrun)
    python bench.py --impl reference [--gpus N ...]            the unmodified reference from baseline/_ref

One JSON line on stdout (rank 0).  `value` is the device-timed whole-job throughput (CUDA events
around exactly K steps, barrier + synchronize on both sides, max over ranks); `e2e` is the same
metric through the public API with the host->device copy of every step's inputs (pinned memory) and
a device->host read of every step's loss inside the timed region.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BATCH = 256  # per GPU (BASELINE.json configs 3 and 4)
RATE = 0.1   # cnn.c:446
MODEL = "mnist-cnn 1x28x28 -> conv3x3s2(16)+relu -> conv3x3s2(32)+relu -> fc200 tanh -> fc200 tanh -> fc10 softmax (360,810 params)"
# Serial cnn.c measured by the survey on the authoring machine: the only published-equivalent number.
BASELINE_IMG_S = 113.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "nccl"])
    ap.add_argument("--algo", default="auto", choices=["auto", "one_shot", "two_shot", "nvls"])
    ap.add_argument("--backend", default="auto", choices=["auto", "simt", "tcgen05"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--dataset", type=int, default=204800, help="device-resident synthetic samples per rank (x784 B; default > L2)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# reference arm
# ------------------------------------------------------------------------------------------------
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
    from mpi_cuda_cnn_b200.utils import write_synthetic_dataset

    # The reference has no flags: 10 epochs, batch 32, fp64, one sample at a time (cnn.c:446-451).
    # K steps of OUR batch would be K*256 samples; the serial reference needs ~9 ms per sample, so it gets
    # a fixed, smaller budget: train_size samples x 10 epochs.
    train_size = 160 * max(world, 1)
    tmp = os.environ.get("B200CNN_REF_TMP") or os.path.join(tempfile.gettempdir(), f"b200cnn_ref_{os.environ.get('MASTER_PORT', '0')}")
    if rank == 0:
        write_synthetic_dataset(tmp, ntrain=train_size, ntest=10)
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
        out = {
            "impl": "reference", "metric": "train_images_per_sec", "value": samples / dt, "unit": "images/s", "n_gpus": args.gpus,
            "steps": samples // 32, "warmup": 0, "ms_per_step": 1e3 * dt / max(samples // 32, 1), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": (samples / dt) / BASELINE_IMG_S, "dtype": "fp64", "data": "synthetic",
            "exit_code": res.returncode,
            "config": {"model": MODEL, "global_batch": 32 if world == 1 else world, "seq_len": None,
                       "parallelism": "none (serial cnn.c)" if world == 1 else f"mpi-dp{world} (cnnmpi.c over an in-repo shared-memory mpi.h shim; no MPI in the image)",
                       "program": os.path.basename(binary), "build": "gcc, reference Makefile flags (no -O)",
                       "samples": samples, "wall_s": dt},
            "note": "The reference's GPU programs (CUDAcnn.cu, CUDAMPI.c/.cu) do not compile (baseline/_ref/BUILD_STATUS.txt); "
                    "this is its unmodified CPU trainer, which uses no GPU. Wall-clock timed (the reference has no timers); "
                    "cnnmpi.c never reads the IDX payload and all-reduces per sample (SURVEY D6-D10), so only its speed is meaningful.",
        }
        print(json.dumps(out))
    return 0


# ------------------------------------------------------------------------------------------------
# this framework
# ------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    from mpi_cuda_cnn_b200 import Engine
    from mpi_cuda_cnn_b200.ops import functional as F200
    from mpi_cuda_cnn_b200.parallel import DataParallelTrainer, init_distributed
    from mpi_cuda_cnn_b200.utils import ClockSampler

    rank, world, device = init_distributed()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device")
    F200.set_backend(args.backend)
    use_graph = not args.no_graph
    mode = "nccl" if args.impl == "nccl" else "fused"
    if world == 1 and mode == "fused":
        engine = Engine(max_batch=BATCH, device=device)
        engine.init_params(seed=0)
        trainer = None
    else:
        trainer = DataParallelTrainer(rank, world, device, batch=BATCH, mode=mode, algo=args.algo)
        engine = trainer.engine

    def step(x, y):
        if trainer is None:
            engine.train_step(x, y, RATE / BATCH, use_graph=use_graph)
        else:
            trainer.train_step(x, y, RATE, use_graph=use_graph)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    stream = torch.cuda.Stream(device=device)
    K, W = args.steps, max(args.warmup, 3)
    n_data = max(args.dataset, BATCH)
    with torch.cuda.stream(stream):
        # Device-resident synthetic shard for this rank (K10), larger than L2 so successive steps read
        # different input lines; each step gathers a random batch from it like the reference's
        # rand() % train_size sampling (cnn.c:455).
        data_x, data_y = F200.synthetic_batch(n_data, seed=1, first_index=rank * n_data, device=device)
        xs = torch.empty((BATCH, 784), dtype=torch.uint8, device=device)
        ys = torch.empty((BATCH,), dtype=torch.uint8, device=device)
        engine.reset_metrics()

        # Data loading overlapped with training: batch i+1 is gathered on a second stream into the other staging
        # buffer while step i runs (the engine keeps one captured graph per staging buffer).
        xs2, ys2 = torch.empty_like(xs), torch.empty_like(ys)
        dstage = [(xs, ys), (xs2, ys2)]
        gather_stream = torch.cuda.Stream(device=device)
        gathered = [torch.cuda.Event(), torch.cuda.Event()]
        freed = [torch.cuda.Event(), torch.cuda.Event()]
        for ev in freed:
            ev.record(stream)

        def device_gather(i):
            with torch.cuda.stream(gather_stream):
                gather_stream.wait_event(freed[i % 2])
                F200.gather_batch(data_x, data_y, BATCH, seed=3, step=i, out=dstage[i % 2])
                gathered[i % 2].record(gather_stream)

        def device_step(i):
            stream.wait_event(gathered[i % 2])
            step(*dstage[i % 2])
            freed[i % 2].record(stream)

        device_gather(0)
        for i in range(W):
            device_gather(i + 1)
            device_step(i)
        stream.synchronize()
        barrier()
        engine.reset_metrics()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(gpu_index=device.index or 0) as clocks:
            ev0.record(stream)
            for i in range(K):
                if i + 1 < K:
                    device_gather(W + i + 1)
                device_step(W + i)
            ev1.record(stream)
            stream.synchronize()
        barrier()
        gather_stream.synchronize()
        ms = ev0.elapsed_time(ev1)
        m_dev = engine.read_metrics()

        # ---- end to end: pinned host inputs -> device, loss read back, every step -------------------
        pool = max(1, min(64, n_data // BATCH))
        host_x = torch.empty((pool, BATCH, 784), dtype=torch.uint8).pin_memory()
        host_y = torch.empty((pool, BATCH), dtype=torch.uint8).pin_memory()
        host_x.copy_(data_x[: pool * BATCH].reshape(pool, BATCH, 784).cpu())
        host_y.copy_(data_y[: pool * BATCH].reshape(pool, BATCH).cpu())
        # Two pinned result slots: the host reads step i-1's loss while step i runs (each step's inputs are still
        # copied from pinned host memory and each step's loss is still read back inside the timed region).
        loss_host = [torch.zeros(4, dtype=torch.float32).pin_memory() for _ in range(2)]
        loss_ready = [torch.cuda.Event(), torch.cuda.Event()]
        metrics = engine.metrics_tensor()
        losses = []

        # Double-buffered device staging: batch i+1 crosses PCIe on the copy stream while batch i trains.
        stage = [(xs, ys), (torch.empty_like(xs), torch.empty_like(ys))]
        copy_stream = torch.cuda.Stream(device=device)
        copied = [torch.cuda.Event(), torch.cuda.Event()]
        consumed = [torch.cuda.Event(), torch.cuda.Event()]
        for ev in consumed:
            ev.record(stream)

        def e2e_upload(i):
            if mode == "fused":
                return  # train_step_host copies on the engine's own copy stream
            sx, sy = stage[i % 2]
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[i % 2])  # the step that last read this buffer has finished
                sx.copy_(host_x[i % pool], non_blocking=True)
                sy.copy_(host_y[i % pool], non_blocking=True)
                copied[i % 2].record(copy_stream)

        def e2e_enqueue(i):
            if mode == "fused":
                # the public host-fed step: H2D of this step's inputs (engine's copy stream, double-buffered), the
                # captured step, D2H of the loss — one native call, nothing blocks
                engine.train_step_host(host_x[i % pool], host_y[i % pool], RATE / BATCH, metrics_out=loss_host[i % 2], stream=stream)
            else:
                sx, sy = stage[i % 2]
                stream.wait_event(copied[i % 2])
                step(sx, sy)
                consumed[i % 2].record(stream)
                loss_host[i % 2].copy_(metrics, non_blocking=True)
            loss_ready[i % 2].record(stream)

        def e2e_collect(i):
            loss_ready[i % 2].synchronize()
            return float(loss_host[i % 2][0])

        for i in range(W):
            e2e_upload(i)
            e2e_enqueue(i)
            e2e_collect(i)
        barrier()
        engine.reset_metrics()
        stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        e2e_upload(0)
        for i in range(K):
            if i + 1 < K:
                e2e_upload(i + 1)
            e2e_enqueue(i)
            if i > 0:
                losses.append(e2e_collect(i - 1))
        losses.append(e2e_collect(K - 1))
        e1.record(stream)
        stream.synchronize()
        barrier()
        ms_e2e = e0.elapsed_time(e1)

    t = torch.tensor([ms, ms_e2e], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = t.tolist()
    if trainer is not None and trainer.comm is not None and trainer.comm.error():
        raise SystemExit("fused communicator reported a peer timeout")
    value = world * BATCH * K / (ms / 1e3)
    e2e = world * BATCH * K / (ms_e2e / 1e3)
    launches = (engine.launches_per_step + 1) * K if mode == "fused" else None
    if rank == 0:
        ck = clocks.summary()
        out = {
            "metric": "train_images_per_sec", "value": value, "unit": "images/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": value / BASELINE_IMG_S,
            "dtype": "tf32" if F200_backend_uses_tc(args.backend) else "fp32", "data": "synthetic", "impl": args.impl,
            "config": {"model": MODEL, "global_batch": BATCH * world, "seq_len": None, "parallelism": f"dp{world}",
                       "per_gpu_batch": BATCH, "update": ("fused p2p allreduce+sgd (%s)" % args.algo) if mode == "fused" else "nccl allreduce + scale/sgd kernel",
                       "cuda_graph": use_graph, "backend": args.backend,
                       "region": (trainer.comm.region.kind if trainer is not None and trainer.comm is not None else None),
                       "nvls_multicast": (trainer.comm.has_multicast if trainer is not None and trainer.comm is not None else None),
                       "l2_policy": f"inputs larger than L2: each step gathers a random batch from a {n_data * 784 / 2**20:.0f} MiB device-resident set",
                       "final_mse_metric": m_dev["mse_sum"] / max(m_dev["count"], 1), "final_accuracy": m_dev["correct"] / max(m_dev["count"], 1)},
            "clocks": {"sm_mhz": ck["sm_mhz"], "sm_max_mhz": ck["sm_max_mhz"], "reasons": ck["reasons"], "samples": ck["samples"]},
            "e2e": {"value": e2e, "unit": "images/s", "ms_per_step": ms_e2e / K, "h2d_bytes_per_step": BATCH * 784 + BATCH,
                    "d2h_bytes_per_step": 16,
                    "last_step_mean_ce": (losses[-1] - losses[-2]) / BATCH if len(losses) > 1 else None},
            "gpu_launches": launches,
            "launches_per_step": engine.launches_per_step + 1,
        }
        print(json.dumps(out))
    if trainer is not None:
        trainer.close()
    else:
        engine.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


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
