#!/bin/bash
# GPU-box job 4 (round 2): 2-rank strong-scaling bench of the C4 snapshot
# (w-plane ownership, NCCL image reduce), as the driver launches it.
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/box2.txt
( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 2 --warmup 3 --second-steps 1 --e2e-steps 1 ) > $O/r02_b_c4_n2.json 2> $O/r02_b_c4_n2.err
cut -c1-400 $O/r02_b_c4_n2.json; tail -8 $O/r02_b_c4_n2.err
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 2 --impl reference --steps 1 --warmup 1 ) > $O/r02_ref_c4_n2.json 2> $O/r02_ref_c4_n2.err
cut -c1-300 $O/r02_ref_c4_n2.json; tail -3 $O/r02_ref_c4_n2.err
