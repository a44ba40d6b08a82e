#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:edge_mlp_aggregate_pair -s 5 -c 3 \
  -o gpurun_out/r02_edge_pair python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_edge_pair.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
