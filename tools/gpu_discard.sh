#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_bf16.py -m gpu -x -q -k "native" 2>&1 | tail -2
timeout 300 python tools/insitu_kernels.py > gpurun_out/insitu_discard.txt 2>&1; tail -36 gpurun_out/insitu_discard.txt
GN_NO_DISCARD=1 timeout 300 python tools/insitu_kernels.py 2>&1 | tail -1
timeout 600 python bench.py --no-cpu-baseline --no-riders --steps 30 > gpurun_out/bench_disc.json 2> gpurun_out/bench_disc.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_disc.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('graph_replay'), d.get('gpu_launches'))"
