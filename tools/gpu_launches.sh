#!/bin/bash
# GPU tests of the 16-bit path, a short bench line, and the in-situ launch list of the C2 step
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_bf16.py tests/test_gpu_fullsize.py tests/test_gpu_ops.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --no-cpu-baseline --no-riders --steps 20 > gpurun_out/q_c2.json 2> gpurun_out/q_c2.err; echo "c2 rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/q_c2.json"))
print("ms/step %.4f value %.0f e2e %.0f graph %s" % (d["ms_per_step"], d["value"], d["e2e"]["value"], (d.get("graph_replay") or {}).get("ms_per_step")))
print("roofline_hbm", {k: (d.get("roofline_hbm") or {}).get(k) for k in ("achieved", "frac", "ms_per_launch")})
PY
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
python tools/summarise_launches.py gpurun_out/launches.csv 2>/dev/null | head -28
