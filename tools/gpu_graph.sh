#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_configs.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/profile_host_infer.py > gpurun_out/host_profile_graph.txt 2>&1; head -3 gpurun_out/host_profile_graph.txt
GN_NO_GRAPH=1 timeout 300 python tools/profile_host_infer.py 2>&1 | head -2
timeout 600 python bench.py --no-cpu-baseline --no-riders --steps 30 > gpurun_out/bench_graph.json 2> gpurun_out/bench_graph.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_graph.err
python -c "
import json; d=json.load(open('gpurun_out/bench_graph.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e'].get('ms_per_step'), d.get('graph_replay'), d.get('gpu_launches'))"
GN_NO_GRAPH=1 timeout 600 python bench.py --no-cpu-baseline --no-riders --steps 30 > gpurun_out/bench_nograph.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/bench_nograph.json'))
print('no graph:', d['value'], d['ms_per_step'], d['e2e']['value'])"
