#!/bin/bash
# 2-GPU job: the single-process unit shards with peer stores, the per-rank path with the
# IPC peer buffers (bench.py under torchrun), and the NCCL fallback.
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/j4_topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_dist.py -m gpu -x -q > gpurun_out/j4_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j4_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/j4_bench2.json 2> gpurun_out/j4_bench2.err; echo "bench2 rc=$?"
SQZ_NO_PEER=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/j4_bench2_nccl.json 2> gpurun_out/j4_bench2_nccl.err; echo "bench2 nccl rc=$?"
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/j4_bench1.json 2> gpurun_out/j4_bench1.err; echo "bench1 rc=$?"
tail -3 gpurun_out/j4_pytest.log; tail -3 gpurun_out/j4_bench2.err; python tools/brief.py < gpurun_out/j4_bench2.json; python tools/brief.py < gpurun_out/j4_bench2_nccl.json; python tools/brief.py < gpurun_out/j4_bench1.json
