#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py -m gpu -x -q -k "tc_linear or fused_linear_pool or native" > gpurun_out/lin_tests.log 2>&1; echo "lin tests rc=$?"; tail -3 gpurun_out/lin_tests.log
export GN_B200_DEBUG_LIB=1 GN_NO_FLUSH=1 GN_TC_DEBUG=1
for mc in 4 1; do for h in 0 24; do
echo "=== MAX_CLUSTER=$mc HACK=$h"
GN_TC_MAX_CLUSTER=$mc GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 704 256 2>&1 | grep -v "^item" | head -20
GN_TC_MAX_CLUSTER=$mc GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 336 1088 2>&1 | head -1
done; done > gpurun_out/lin_stages2.txt 2>&1
cat gpurun_out/lin_stages2.txt
unset GN_B200_DEBUG_LIB GN_NO_FLUSH GN_TC_DEBUG
timeout 300 python tools/insitu_kernels.py > gpurun_out/insitu_c2_c.txt 2>&1; tail -32 gpurun_out/insitu_c2_c.txt | grep -i "tc_linear\|sum of"
GN_TC_MAX_CLUSTER=1 timeout 300 python tools/insitu_kernels.py > gpurun_out/insitu_c2_d.txt 2>&1; tail -32 gpurun_out/insitu_c2_d.txt | grep -i "tc_linear\|sum of"
timeout 600 python bench.py --no-cpu-baseline --no-riders --steps 30 > gpurun_out/bench_lin.json 2> gpurun_out/bench_lin.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_lin.err
python -c "
import json; d=json.load(open('gpurun_out/bench_lin.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('graph_replay'), d.get('gpu_launches'))"
