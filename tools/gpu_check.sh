#!/bin/bash
# One B200 box: GPU parity tests, smoke, the default bench line (riders included), reference arm.
# Everything lands in gpurun_out/ (merged back by gpurun); summaries are copied to profiles/ by hand.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "no-alloc alone rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_n1.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n1.json"))
print("value %.0f ev/s  %.4f ms/step  e2e %.0f  graph %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], (d.get("graph_replay") or {}).get("ms_per_step")))
print("roofline", {k: d["roofline"].get(k) for k in ("achieved", "frac", "ms_per_launch", "achieved_unfactored")})
print("roofline_hbm", {k: (d.get("roofline_hbm") or {}).get(k) for k in ("achieved", "frac", "ms_per_launch")})
for k in ("fp32", "c4", "sharded", "train"):
    v = d.get(k) or {}
    print(k, {kk: v.get(kk) for kk in ("value", "ms_per_step", "error", "edge_kernel", "events_per_rank")})
for k, v in (d.get("c5") or {}).items():
    print("c5", k, {kk: v.get(kk) for kk in ("value", "ms_per_step", "nodes_per_s", "error")})
print("cpu_baseline", d.get("cpu_baseline"))
PY
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 120 python tools/time_edge.py c2 2>/dev/null; timeout 120 python tools/time_edge.py c4 2>/dev/null
