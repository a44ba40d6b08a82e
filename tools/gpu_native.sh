#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py -m gpu -x -q -k "native" > gpurun_out/native_tests.log 2>&1; echo "native tests rc=$?"; tail -15 gpurun_out/native_tests.log
timeout 300 python tools/profile_host_infer.py > gpurun_out/host_profile_native.txt 2>&1; head -4 gpurun_out/host_profile_native.txt
timeout 600 python bench.py --no-cpu-baseline --no-riders --steps 30 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_native.err
python -c "
import json; d=json.load(open('gpurun_out/bench_native.json'))
print(d['value'], d['ms_per_step'], d['e2e'], d.get('graph_replay'), d.get('gpu_launches'))"
