#!/bin/bash
# End-of-round record on one B200 box: GPU tests, no-alloc test alone, smoke, the default bench line
# (riders + CPU baseline), the reference arm, in-step kernel times, host profile.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "no-alloc alone rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 1200 python bench.py > gpurun_out/r02_final_bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_n1.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/r02_final_bench_n1.json"))
print("value %.0f ev/s  %.4f ms/step  e2e %.0f (%.4f ms)  e2e_dg %.0f  graph %s  launches %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e_device_graph"]["value"], (d.get("graph_replay") or {}).get("ms_per_step"), d.get("gpu_launches")))
print("roofline", {k: d["roofline"].get(k) for k in ("achieved", "frac", "ms_per_launch", "achieved_unfactored")})
print("roofline_hbm", {k: (d.get("roofline_hbm") or {}).get(k) for k in ("achieved", "frac", "ms_per_launch")})
for k in ("fp32", "c4", "sharded", "train"):
    v = d.get(k) or {}
    print(k, {kk: v.get(kk) for kk in ("value", "ms_per_step", "error", "edge_kernel", "events_per_rank")})
for k, v in (d.get("c5") or {}).items():
    print("c5", k, {kk: v.get(kk) for kk in ("value", "ms_per_step", "nodes_per_s", "error")})
print("cpu_baseline", d.get("cpu_baseline"))
print("clocks", d.get("clocks"))
PY
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_final_bench_reference_arm.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/r02_final_bench_reference_arm.json | cut -c1-300
timeout 300 python tools/insitu_kernels.py > gpurun_out/r02_insitu_c2.txt 2>&1; tail -3 gpurun_out/r02_insitu_c2.txt
timeout 300 python tools/insitu_kernels.py 256 500 bf16 c4 > gpurun_out/r02_insitu_c4.txt 2>&1; tail -1 gpurun_out/r02_insitu_c4.txt
timeout 300 python tools/insitu_kernels.py 1024 60 fp32 > gpurun_out/r02_insitu_fp32.txt 2>&1; tail -1 gpurun_out/r02_insitu_fp32.txt
timeout 300 python tools/profile_host_infer.py > gpurun_out/r02_host_profile_infer.txt 2>&1; head -2 gpurun_out/r02_host_profile_infer.txt
