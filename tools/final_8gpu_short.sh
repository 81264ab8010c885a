# Eight GPUs, last refresh of the round: N = 8 (K = 20 and 200) and N = 4 (K = 20) with the final code.
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $2 "${@:3}"; }
run 8 29631 bench.py --gpus 8 --steps 20 --warmup 3 --parity-steps 0 > gpurun_out/final_bench_n8_k20.json 2> gpurun_out/final_bench_n8_k20.err
run 8 29632 bench.py --gpus 8 --steps 200 --warmup 20 --parity-steps 0 > gpurun_out/final_bench_n8_k200.json 2> gpurun_out/final_bench_n8_k200.err
run 4 29633 bench.py --gpus 4 --steps 20 --warmup 3 --parity-steps 0 > gpurun_out/final_bench_n4_k20.json 2> gpurun_out/final_bench_n4_k20.err
for f in final_bench_n8_k20 final_bench_n8_k200 final_bench_n4_k20; do python -c "
import json;d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1]);print('$f',round(d['value']),d['ms_per_step'],d['e2e']['ms_per_step'],d.get('baselines',{}).get('nccl_ms_per_step'),d['clocks'])" || tail -5 gpurun_out/$f.err; done
