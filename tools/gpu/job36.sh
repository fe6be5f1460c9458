#!/bin/bash
# 8-GPU job with the final kernels: multi-GPU tests, strong scaling of config 3 at N = 8, 4, 2, cohort (config 5).
mkdir -p gpurun_out
run() { # N tag [env...]
  n=$1; tag=$2; shift 2
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/j36_$tag.json 2> gpurun_out/j36_$tag.err
  echo "$tag rc=$?"; python tools/brief.py < gpurun_out/j36_$tag.json
}
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_dist.py -m gpu -x -q > gpurun_out/j36_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j36_pytest.log; tail -2 gpurun_out/j36_pytest.log
run 8 n8 A=1
run 4 n4 A=1
run 2 n2 A=1
timeout 300 python bench.py --steps 10 --warmup 3 --workload config5 --gpus 8 > gpurun_out/j36_config5.json 2> gpurun_out/j36_config5.err; head -c 600 gpurun_out/j36_config5.json
