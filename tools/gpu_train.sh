#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"
tail -3 gpurun_out/pytest_gpu.log
python tools/profile_host.py 2>&1 | cut -c1-150 > gpurun_out/profile_host.txt; head -4 gpurun_out/profile_host.txt
for i in 1 2; do timeout 300 python bench.py --mode train --no-cpu-baseline > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; echo "train rc=$?"; cut -c1-200 gpurun_out/bench_train.json; echo; done
timeout 300 python bench.py --no-cpu-baseline --mode infer > gpurun_out/bench_q_c2.json 2> gpurun_out/bench_q_c2.err; echo "c2 rc=$?"; cut -c1-200 gpurun_out/bench_q_c2.json; echo
if [ -n "$1" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv --log-file gpurun_out/train_launches.csv \
  python bench.py --mode train --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_train.log 2>&1; echo "ncu train rc=$?"
fi
