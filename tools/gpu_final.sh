#!/bin/bash
# what the round driver runs at the end: GPU tests, the no-alloc test alone, smoke, both bench arms
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "no-alloc alone rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 300 python bench.py --impl reference --gpus 1 --steps 5 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$? lines=$(wc -l < gpurun_out/bench_ref.json)"
timeout 600 python bench.py --gpus 1 --steps 30 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$? lines=$(wc -l < gpurun_out/bench_n1.json)"
cut -c1-220 gpurun_out/bench_n1.json
