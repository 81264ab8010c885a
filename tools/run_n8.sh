set -x
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
run 8 29601 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2g_bench_n8_k20.json 2> gpurun_out/r2g_bench_n8_k20.err
run 8 29602 bench.py --gpus 8 --steps 200 --warmup 20 --parity-steps 0 > gpurun_out/r2g_bench_n8_k200.json 2> gpurun_out/r2g_bench_n8_k200.err
run 4 29603 bench.py --gpus 4 --steps 20 --warmup 5 --parity-steps 0 > gpurun_out/r2g_bench_n4_k20.json 2> gpurun_out/r2g_bench_n4_k20.err
run 4 29604 bench.py --gpus 4 --steps 200 --warmup 20 --parity-steps 0 > gpurun_out/r2g_bench_n4_k200.json 2> gpurun_out/r2g_bench_n4_k200.err
run 8 29605 bench/allreduce_sweep.py --max-bytes 67108864 --out gpurun_out/r2g_sweep_w8.json > gpurun_out/r2g_sweep_w8.log 2>&1
timeout 400 python -m pytest tests/test_gpu_comm.py -q -m gpu -x -k "multi_gpu" > gpurun_out/r2g_pytest_n8.log 2>&1
tail -3 gpurun_out/r2g_pytest_n8.log
TIMELINE_OUT=gpurun_out/r2g_timeline_n8 run 8 29606 tools/timeline.py > /dev/null 2> gpurun_out/r2g_timeline.err
python - <<PY
import json
for f in ["n8_k20","n8_k200","n4_k20","n4_k200"]:
    try:
        d=json.loads(open("gpurun_out/r2g_bench_%s.json"%f).read().strip().splitlines()[-1]); print(f, round(d["ms_per_step"]*1e3,1), round(d["e2e"]["ms_per_step"]*1e3,1), d.get("baselines",{}).get("nccl_ms_per_step"), d["config"]["nvls_multicast"])
    except Exception as ex: print(f, "ERR", ex)
try:
    d=json.load(open("gpurun_out/r2g_sweep_w8.json"))
    for b in sorted({r["bytes"] for r in d}):
        print(b, " ".join("%s=%.1f"%(r["path"].split("/")[-1][:9]+("s" if "sgd" in r["path"] else ""), r["us"]) for r in d if r["bytes"]==b and ("sgd" in r["path"] or r["path"]=="nccl")))
except Exception as ex: print("sweep ERR", ex)
PY
