#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py -m gpu -x -q -k "tc_linear or fused_linear_pool" > gpurun_out/lin_tests.log 2>&1; echo "lin tests rc=$?"; tail -5 gpurun_out/lin_tests.log
timeout 900 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_training.py tests/test_gpu_backward_ops.py -m gpu -x -q > gpurun_out/lin_tests2.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/lin_tests2.log
export GN_B200_DEBUG_LIB=1
for h in 0 24; do
  echo "== HACK=$h warm, cluster / no cluster"
  GN_NO_FLUSH=1 GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 704 256 2>&1 | head -1
  GN_NO_FLUSH=1 GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 336 1088 2>&1 | head -1
  GN_TC_MAX_CLUSTER=1 GN_NO_FLUSH=1 GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 704 256 2>&1 | head -1
  GN_TC_MAX_CLUSTER=1 GN_NO_FLUSH=1 GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 336 1088 2>&1 | head -1
done 2>&1 | tee gpurun_out/lin_ablate5.txt
GN_NO_FLUSH=1 GN_TC_DEBUG=1 timeout 120 python tools/trace_linear.py 61443 704 256 > gpurun_out/lin_trace4.txt 2>&1; head -8 gpurun_out/lin_trace4.txt
unset GN_B200_DEBUG_LIB
timeout 300 python tools/insitu_kernels.py > gpurun_out/insitu_c2_b.txt 2>&1; tail -32 gpurun_out/insitu_c2_b.txt
timeout 600 python bench.py --no-cpu-baseline --no-riders --steps 30 > gpurun_out/bench_lin.json 2> gpurun_out/bench_lin.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_lin.err
python -c "
import json; d=json.load(open('gpurun_out/bench_lin.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('graph_replay'), d.get('gpu_launches'))"
