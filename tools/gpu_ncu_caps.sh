#!/bin/bash
# ncu --set full captures of the three heaviest kernels of the C2 step (cold cache, serialised)
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:edge_mlp_aggregate_pair -s 5 -c 3 \
  -o gpurun_out/r02_edge_pair -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_edge_pair.log 2>&1; echo "ncu edge rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_linear_kernel -s 12 -c 6 \
  -o gpurun_out/r02_tc_linear -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_tc_linear.log 2>&1; echo "ncu linear rc=$?"
ls -la gpurun_out/*.ncu-rep
