# One GPU, end of round: the whole GPU test-suite, the bench lines the driver asks for, timelines, launch table and ONE ncu capture of the
# pooled topology's new kernels.  Everything lands in gpurun_out/ (copied to profiles/ afterwards).
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python -m pytest tests -q -m gpu > gpurun_out/pytest_gpu_n1.log 2>&1; tail -3 gpurun_out/pytest_gpu_n1.log
python bench.py --steps 20 --warmup 3 > gpurun_out/final_bench_n1_k20.json 2> gpurun_out/final_bench_n1_k20.err
python bench.py --steps 200 --warmup 20 > gpurun_out/final_bench_n1_k200.json 2> gpurun_out/final_bench_n1_k200.err
python bench.py --pool --steps 200 --warmup 20 > gpurun_out/final_bench_pool_n1_k200.json 2> gpurun_out/final_bench_pool.err
python bench.py --impl reference > gpurun_out/final_bench_reference_n1.json 2>&1
python tools/timeline.py 256 40 > gpurun_out/timeline_step_1gpu.txt 2>/dev/null
python tools/timeline.py 256 40 --pool > gpurun_out/timeline_step_pool_1gpu.txt 2>/dev/null
python tools/bench_kernels.py > gpurun_out/per_op_times_in_graph.txt 2>&1
python tools/diag_engine.py 256 > gpurun_out/accuracy_vs_fp64.txt 2>&1
python tools/diag_engine.py 256 --pool > gpurun_out/accuracy_vs_fp64_pool.txt 2>&1
CMD="python bench.py --steps 4 --warmup 3 --no-baseline --parity-steps 0 --dataset 4096"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 40 --csv --log-file gpurun_out/launches_step_1gpu.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD --pool > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"halo|bwd_weight_kernel|conv1_pool|maxpool_nhwc" -s 24 -c 8 -f -o gpurun_out/prof_pool_r2 $CMD --pool > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out/prof_pool_r2.ncu-rep; tail -2 gpurun_out/ncu2.log
for f in final_bench_n1_k20 final_bench_n1_k200 final_bench_pool_n1_k200; do python -c "
import json;d=json.loads(open('gpurun_out/$f.json').read().strip().splitlines()[-1]);print('$f',round(d['value']),d['ms_per_step'],d['e2e']['ms_per_step'],d['gpu_launches'],d['parity']['ok'],d['clocks'])"; done
cut -c1-400 gpurun_out/final_bench_reference_n1.json
