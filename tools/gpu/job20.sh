#!/bin/bash
# Full GPU suite + bench + the ncu evidence for profiles/r02_*: launch lists (config 3 and 2),
# full captures of the two dominant kernels, and the executed fp64 instruction counters.
mkdir -p gpurun_out
git_head=$(cat .git_head 2>/dev/null)
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/j20_bench.json 2> gpurun_out/j20_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/j20_bench_ref.json 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/j20_launches_c3.csv python bench.py --steps 2 --warmup 3 > gpurun_out/j20_ncu_l3.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/j20_launches_c2.csv python bench.py --steps 5 --warmup 3 --workload config2 > gpurun_out/j20_ncu_l2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:baf_grid_colwarp -s 2 -c 1 -o gpurun_out/j20_c3_colwarp -f python tools/quick_c3.py > gpurun_out/j20_ncu_f3.log 2>&1
WORKLOAD=c2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:baf_grid_colwarp -s 2 -c 1 -o gpurun_out/j20_c2_colwarp -f python tools/quick_c3.py > gpurun_out/j20_ncu_f2.log 2>&1
M=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum,gpu__time_duration.sum
timeout 900 ncu --metrics $M --clock-control none -k regex:"baf_grid_colwarp|baf_colterm" -s 4 -c 2 --csv --log-file gpurun_out/j20_flops_c3.csv python tools/quick_c3.py > gpurun_out/j20_ncu_m3.log 2>&1
WORKLOAD=c2 timeout 900 ncu --metrics $M --clock-control none -k regex:"baf_grid_colwarp|baf_colterm" -s 4 -c 2 --csv --log-file gpurun_out/j20_flops_c2.csv python tools/quick_c3.py > gpurun_out/j20_ncu_m2.log 2>&1
python tools/brief.py < gpurun_out/j20_bench.json; tail -2 gpurun_out/j20_bench.err
