#!/bin/bash
# Round-2 end state: launch lists (bf16 C2, fp32 C2) and ncu --set full captures of the fp32 edge kernel
# and of the tcgen05 GEMM launches of the bf16 step.  Each capture only after the plain command exited 0.
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/p_c2.json 2> gpurun_out/p_c2.err; echo "c2 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_final_launches_bf16.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_l1.log 2>&1; echo "ncu launches bf16 rc=$?"
timeout 300 python bench.py --precision fp32 --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/p_fp32.json 2> gpurun_out/p_fp32.err; echo "fp32 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 500 --csv --log-file gpurun_out/r02_final_launches_fp32.csv \
  python bench.py --precision fp32 --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_l2.log 2>&1; echo "ncu launches fp32 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:edge_mlp_aggregate_x3 -s 6 -c 2 \
  -o gpurun_out/r02_edge_x3 -f python bench.py --precision fp32 --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_x3.log 2>&1; echo "ncu x3 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_linear_kernel -s 14 -c 7 \
  -o gpurun_out/r02_tc_linear_final -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_lin.log 2>&1; echo "ncu linear rc=$?"
ls -la gpurun_out/*.ncu-rep
