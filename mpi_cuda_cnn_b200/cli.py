"""The four entrypoints of the reference, same command line (four positional IDX paths).

    python -m mpi_cuda_cnn_b200.cli cnn     [flags] train-images train-labels test-images test-labels   (cnn.c)
    python -m mpi_cuda_cnn_b200.cli cnnmpi  -np N [flags] ...                                            (cnnmpi.c)
    python -m mpi_cuda_cnn_b200.cli cudacnn [flags] ...                                                  (CUDAcnn.cu)
    python -m mpi_cuda_cnn_b200.cli cudampi -np N [flags] ...                                            (CUDAMPI.c/.cu)

cnn / cnnmpi run the native CPU binaries.  cudacnn / cudampi train on the GPU through the native engine;
hyper-parameter flags default to the reference's literals (/root/reference/cnn.c:413, :446-449), the GPU
programs default to the benchmark's batch of 256 per GPU.  stderr lines and exit codes (100 usage,
111 file error) follow the reference (cnn.c:412, :432-518; cnnmpi.c:459, :468-470).
"""
import argparse
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
E_USAGE, E_FILE = 100, 111


def _bin(name):
    path = os.path.join(ROOT, "bin", name)
    if not os.path.exists(path):
        from . import _native

        _native.build(("cpu",))
    return path


def _gpu_parser(dp):
    ap = argparse.ArgumentParser(add_help=True)
    ap.add_argument("paths", nargs="*")
    ap.add_argument("--rate", type=float, default=0.1)
    ap.add_argument("--nepoch", type=int, default=10)
    ap.add_argument("--batch-size", type=int, default=256)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--log-every", type=int, default=1000)
    ap.add_argument("--synthetic", type=int, default=0)
    ap.add_argument("--synthetic-test", type=int, default=0)
    ap.add_argument("--pool", action="store_true")
    ap.add_argument("--backend", default="auto", choices=["auto", "simt", "tcgen05"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--save")
    ap.add_argument("--load")
    ap.add_argument("--timing", action="store_true")
    if dp:
        ap.add_argument("-np", "--np", type=int, default=0)
        ap.add_argument("--update", default="fused", choices=["fused", "nccl"])
        ap.add_argument("--algo", default="auto", choices=["auto", "one_shot", "two_shot", "nvls", "ll_one_shot", "ll_two_shot"])
    return ap


def _load_data(args):
    import numpy as np

    from .utils import make_synthetic, read_idx

    if args.synthetic > 0 and len(args.paths) < 4:
        ntest = args.synthetic_test or max(100, args.synthetic // 5)
        return make_synthetic(args.synthetic, seed=1) + make_synthetic(ntest, seed=2)
    try:
        tx, ty, ex, ey = (read_idx(p) for p in args.paths[:4])
    except OSError as exc:
        print(exc, file=sys.stderr)
        return None
    if tx.ndim != 3 or len(tx) != len(ty) or ex.ndim != 3 or len(ex) != len(ey):
        print("image / label files do not match", file=sys.stderr)
        return None
    return tx, ty, ex, ey


def _evaluate(engine, ex, ey, batch, device):
    import torch

    x = torch.from_numpy(ex.reshape(len(ex), -1)).to(device)
    y = torch.from_numpy(ey).to(device)
    ncorrect = 0
    for i in range(0, len(ex), batch):
        engine.reset_metrics()
        engine.forward(x[i:i + batch].contiguous(), y[i:i + batch].contiguous())
        ncorrect += int(engine.read_metrics()["correct"])
        if (i % 1000) < batch:
            print(f"i={i}", file=sys.stderr)
    return ncorrect


def _train_gpu(args, rank=0, world=1, device=None, trainer=None):
    """Shared loop of cudacnn (world == 1) and cudampi."""
    import numpy as np
    import torch

    from . import Engine, mnist_layers, pool_layers
    from .ops import functional as F200

    data = _load_data(args)
    if data is None:
        return E_FILE
    tx, ty, ex, ey = data
    F200.set_backend(args.backend)
    device = device or torch.device("cuda:0")
    B = args.batch_size
    if trainer is None:
        engine = Engine(pool_layers() if args.pool else mnist_layers(), max_batch=B, device=device)
        engine.init_params(seed=args.seed)
    else:
        engine = trainer.engine
    if args.load:
        # Rank 0 reads the checkpoint, then EVERY replica gets a copy (the trainer's constructor has already
        # broadcast the seed initialisation; loading on rank 0 alone would leave the other ranks on it).
        if rank == 0:
            engine.set_params(np.fromfile(args.load, dtype=np.float32))
        if trainer is not None:
            trainer.sync_params(0)
    train_size = len(tx)
    shard = train_size // world  # contiguous shard per rank, remainder dropped (cnnmpi.c:457-458)
    start, end = shard * rank, shard * (rank + 1)
    if world > 1:
        print(f"{rank} {start} {end}", file=sys.stderr)
    if rank == 0:
        print("training...", file=sys.stderr)
    stream = torch.cuda.Stream(device=device)
    with torch.cuda.stream(stream):
        dx = torch.from_numpy(tx.reshape(train_size, -1)).to(device)
        dy = torch.from_numpy(ty).to(device)
        xs = torch.empty((B, dx.shape[1]), dtype=torch.uint8, device=device)
        ys = torch.empty((B,), dtype=torch.uint8, device=device)
        steps_per_epoch = max(shard // B, 1)
        engine.reset_metrics()
        seen, next_log, step = 0, 0, 0
        t0 = time.perf_counter()
        for epoch in range(args.nepoch):
            if world > 1 and rank == 0:
                print(f"epoch = {epoch}", file=sys.stderr)
            for _ in range(steps_per_epoch):
                # random mini-batch with replacement from this rank's shard (cnn.c:455), gathered on the device
                F200.gather_batch(dx, dy, B, seed=args.seed + 1, step=step * world + rank, shard_begin=start, shard_len=shard, out=(xs, ys))
                if trainer is None:
                    engine.train_step(xs, ys, args.rate / B, use_graph=not args.no_graph)
                else:
                    trainer.train_step(xs, ys, args.rate, use_graph=not args.no_graph)
                step += 1
                seen += B
                if seen > next_log:
                    if trainer is not None:
                        trainer.check_comm()  # a peer that never showed up must fail the run, not corrupt it silently
                    if rank == 0:
                        m = engine.read_metrics()
                        engine.reset_metrics()
                        if m["count"] > 0:
                            print(f"i={seen - B}, error={m['mse_sum'] / m['count']:.4f}", file=sys.stderr)
                    next_log += max(args.log_every, B)
        stream.synchronize()
        if trainer is not None:
            trainer.check_comm()
        dt = time.perf_counter() - t0
        if rank == 0 and args.timing and step:
            print(f"train_images_per_sec={step * B * world / dt:.1f} ({step * B * world} samples, {world} GPU(s), {dt:.3f} s)", file=sys.stderr)
        rc = 0
        if rank == 0:
            if args.save:
                engine.get_params().tofile(args.save)
            print("testing...", file=sys.stderr)
            ncorrect = _evaluate(engine, ex, ey, B, device)
            print(f"ntests={len(ex)}, ncorrect={ncorrect}", file=sys.stderr)
    return rc


def cudacnn(argv):
    args = _gpu_parser(False).parse_args(argv)
    if args.synthetic <= 0 and len(args.paths) < 4:
        return E_USAGE
    return _train_gpu(args)


def _cudampi_rank(rank, world, argv):
    import torch
    import torch.distributed as dist

    from . import mnist_layers, pool_layers
    from .parallel import DataParallelTrainer, init_distributed

    args = _gpu_parser(True).parse_args(argv)
    rank, world, device = init_distributed()
    if rank == 0:
        from .parallel import topology

        topo = topology()
        print(f"topology: {topo['gpus']} GPUs, NVLinks per GPU {topo['nvlinks']}, all peers reachable: "
              f"{all(all(r) for r in topo['peer'])}", file=sys.stderr)
    trainer = DataParallelTrainer(rank, world, device, layers=pool_layers() if args.pool else mnist_layers(),
                                  batch=args.batch_size, mode=args.update, seed=args.seed, algo=args.algo)
    rc = _train_gpu(args, rank, world, device, trainer)
    print(f"rank {rank} param_checksum={trainer.param_checksum():.17g}", file=sys.stderr)
    if world > 1:
        dist.barrier()
    trainer.close()
    if rc:
        raise SystemExit(rc)


def cudampi(argv):
    args = _gpu_parser(True).parse_args(argv)
    if args.synthetic <= 0 and len(args.paths) < 4:
        return E_USAGE
    if args.synthetic <= 0 and not all(os.path.exists(p) for p in args.paths[:4]):
        print("cannot open input files", file=sys.stderr)
        return E_FILE
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ and args.np <= 0:
        _cudampi_rank(0, 1, argv)  # rank / world come from the environment (torchrun)
        return 0
    import torch

    from .parallel import spawn

    world = args.np or torch.cuda.device_count()
    spawn(_cudampi_rank, world, args=(argv,))
    return 0


def main(argv=None):
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv or argv[0] in ("-h", "--help"):
        print(__doc__, file=sys.stderr)
        return E_USAGE
    prog, rest = argv[0], argv[1:]
    if prog in ("cnn", "cnnmpi"):
        return subprocess.call([_bin(prog)] + rest)
    if prog == "cudacnn":
        return cudacnn(rest)
    if prog == "cudampi":
        return cudampi(rest)
    print(f"unknown program {prog!r}; expected cnn, cnnmpi, cudacnn or cudampi", file=sys.stderr)
    return E_USAGE


if __name__ == "__main__":
    sys.exit(main())
