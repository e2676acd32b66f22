#!/bin/bash
# bench at N GPUs (torchrun) + the 2-rank GPU test
N=${1:-2}; T=${2:-r02f}; O=gpurun_out; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --frames ${FRAMES:-default} > $O/${T}_bench_n$N.json 2> $O/${T}_bench_n$N.err; echo "bench rc=$?"; tail -5 $O/${T}_bench_n$N.err
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -2
