#!/bin/bash
# quick loop: all GPU tests, C2/C4 inference lines, kNN microbench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python tools/bench_knn.py > gpurun_out/bench_knn.jsonl 2> gpurun_out/bench_knn.err; echo "bench_knn rc=$?"
cat gpurun_out/bench_knn.jsonl
timeout 300 python bench.py --no-cpu-baseline --mode infer > gpurun_out/bench_q_c2.json 2> gpurun_out/bench_q_c2.err; echo "c2 rc=$?"
timeout 300 python bench.py --workload c4 --events 256 --pulses 500 --no-cpu-baseline --mode infer > gpurun_out/bench_q_c4.json 2> gpurun_out/bench_q_c4.err; echo "c4 rc=$?"
cut -c1-330 gpurun_out/bench_q_c2.json; echo; cut -c1-330 gpurun_out/bench_q_c4.json
if [ -n "$1" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --mode infer > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
fi
