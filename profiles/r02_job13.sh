#!/bin/bash
# GPU-box job 13 (round 2): strong scaling of the C4 snapshot at N GPUs, as the
# driver launches it (torchrun, one rank per GPU). usage: r02_job13.sh N
N=${1:-8}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8 > $O/box_n$N.txt
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 2 --warmup 3 --second-steps 1 --e2e-steps 1 ) > $O/r02_bench_C4_n$N.json 2> $O/r02_bench_C4_n$N.err
cut -c1-300 $O/r02_bench_C4_n$N.json; tail -6 $O/r02_bench_C4_n$N.err
