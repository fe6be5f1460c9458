#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/j31_ref2.json 2> gpurun_out/j31_ref2.err; echo "ref N=2 rc=$?"; head -c 500 gpurun_out/j31_ref2.json; echo; tail -2 gpurun_out/j31_ref2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 > gpurun_out/j31_n2.json 2> gpurun_out/j31_n2.err; echo "ours N=2 rc=$?"; python tools/brief.py < gpurun_out/j31_n2.json
