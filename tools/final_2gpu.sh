# Two GPUs, end of round: the whole GPU test-suite (multi-GPU tests included) and the data-parallel bench lines.
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu_n2.log 2>&1; tail -4 gpurun_out/pytest_gpu_n2.log
run 2 29611 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/final_bench_n2_k20.json 2> gpurun_out/final_bench_n2_k20.err
run 2 29612 bench.py --gpus 2 --steps 200 --warmup 20 --parity-steps 0 > gpurun_out/final_bench_n2_k200.json 2> gpurun_out/final_bench_n2_k200.err
run 2 29613 bench.py --gpus 2 --pool --steps 200 --warmup 20 --parity-steps 0 > gpurun_out/final_bench_pool_n2_k200.json 2> gpurun_out/final_bench_pool_n2.err
for f in final_bench_n2_k20 final_bench_n2_k200 final_bench_pool_n2_k200; do python -c "
import json;d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1]);print('$f',round(d['value']),d['ms_per_step'],d['e2e']['ms_per_step'],d.get('baselines',{}).get('nccl_ms_per_step'),d['clocks'])" || tail -5 gpurun_out/$f.err; done
