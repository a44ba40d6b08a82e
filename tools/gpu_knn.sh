#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_knn.py tests/test_gpu_fullsize.py -x -q > gpurun_out/pytest_knn.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_knn.log
timeout 300 python tools/bench_knn.py > gpurun_out/bench_knn.jsonl 2> gpurun_out/bench_knn.err; echo "bench_knn rc=$?"
cat gpurun_out/bench_knn.jsonl
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest all rc=$?"
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --no-cpu-baseline --mode infer > gpurun_out/bench_knn_c2.json 2> gpurun_out/bench_knn_c2.err; echo "c2 rc=$?"
timeout 300 python bench.py --workload c4 --events 256 --pulses 500 --no-cpu-baseline --mode infer > gpurun_out/bench_knn_c4.json 2> gpurun_out/bench_knn_c4.err; echo "c4 rc=$?"
cut -c1-330 gpurun_out/bench_knn_c2.json; echo; cut -c1-330 gpurun_out/bench_knn_c4.json
