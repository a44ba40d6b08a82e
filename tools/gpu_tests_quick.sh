#!/bin/bash
# all GPU tests + the no-alloc test alone + smoke + a short bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "no-alloc alone rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --no-cpu-baseline --no-riders --steps 30 > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_quick.json'))
print(d['value'], d['ms_per_step'], d['e2e']['value'], d.get('graph_replay'), d.get('gpu_launches'))"
