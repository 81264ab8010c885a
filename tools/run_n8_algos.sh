run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 --steps 100 --warmup 10 --parity-steps 0 --no-baseline "${@:2}" 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step']*1e3,1), round(d['e2e']['ms_per_step']*1e3,1))"; }
echo "auto: $(run 29611)"
echo "nvls: $(run 29612 --algo nvls)"
echo "two_shot: $(run 29613 --algo two_shot)"
echo "auto blocks16: $(B200CNN_COMM_BLOCKS=16 run 29614)"
echo "auto blocks64: $(B200CNN_COMM_BLOCKS=64 run 29615)"
echo "auto nofusedtail: $(B200CNN_FUSED_TAIL=0 run 29616)"
