# Eight GPUs, end of round: data-parallel bench lines at N = 8 and 4 (driver's K = 20 and the steadier K = 200) and the multi-GPU tests.
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
run 8 29621 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/final_bench_n8_k20.json 2> gpurun_out/final_bench_n8_k20.err
run 8 29622 bench.py --gpus 8 --steps 200 --warmup 20 --parity-steps 0 > gpurun_out/final_bench_n8_k200.json 2> gpurun_out/final_bench_n8_k200.err
run 4 29623 bench.py --gpus 4 --steps 20 --warmup 3 --parity-steps 0 > gpurun_out/final_bench_n4_k20.json 2> gpurun_out/final_bench_n4_k20.err
run 4 29624 bench.py --gpus 4 --steps 200 --warmup 20 --parity-steps 0 > gpurun_out/final_bench_n4_k200.json 2> gpurun_out/final_bench_n4_k200.err
run 8 29625 bench.py --gpus 8 --pool --steps 200 --warmup 20 --parity-steps 0 > gpurun_out/final_bench_pool_n8_k200.json 2> gpurun_out/final_bench_pool_n8.err
timeout 500 python -m pytest tests/test_gpu_comm.py -q -m gpu -k "multi_gpu and (auto or pooled or nvls or nccl)" > gpurun_out/pytest_gpu_n8.log 2>&1; tail -3 gpurun_out/pytest_gpu_n8.log
for f in final_bench_n8_k20 final_bench_n8_k200 final_bench_n4_k20 final_bench_n4_k200 final_bench_pool_n8_k200; do python -c "
import json;d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1]);print('$f',round(d['value']),d['ms_per_step'],d['e2e']['ms_per_step'],d.get('baselines',{}).get('nccl_ms_per_step'),d['config']['nvls_multicast'],d['clocks'])" || tail -5 gpurun_out/$f.err; done
