# One GPU: refresh the single-GPU evidence under gpurun_out/ (copied to profiles/ afterwards).
CMD="python bench.py --steps 4 --warmup 3 --no-baseline --parity-steps 0 --dataset 4096"
python tools/timeline.py > gpurun_out/timeline_step_1gpu.txt 2>/dev/null
python tools/bench_kernels.py > gpurun_out/per_op_times_in_graph.txt 2>&1
python tools/diag_chain.py 256 > gpurun_out/diag_chain_m256.txt 2>&1
python bench.py --steps 20 --warmup 5 > gpurun_out/final_bench_n1_k20.json 2> gpurun_out/final_bench_n1_k20.err
python bench.py --steps 200 --warmup 20 > gpurun_out/final_bench_n1_k200.json 2> gpurun_out/final_bench_n1_k200.err
python bench.py --impl reference > gpurun_out/final_bench_reference_n1.json 2>&1
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 40 --csv --log-file gpurun_out/launches_step_1gpu.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"tc_conv1_bwd_weight|tc_conv3x3|tc_gemm_kernel" -s 30 -c 12 -f -o gpurun_out/prof_r2 $CMD > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out/prof_r2.ncu-rep; tail -2 gpurun_out/ncu2.log; cut -c1-300 gpurun_out/final_bench_n1_k20.json
