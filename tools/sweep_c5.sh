#!/bin/bash
# BASELINE configs[4]: pulses/event sweep 16..2048, direction + energy heads, ~1 M nodes per GPU.
#   bash tools/sweep_c5.sh [out.jsonl]        (one bench.py line per point, inputs resident + e2e)
out=${1:-gpurun_out/sweep_c5.jsonl}
mkdir -p "$(dirname "$out")"; : > "$out"
for n in 16 32 64 128 256 512 1024 2048; do
  ev=$((1048576 / n))
  timeout 600 python bench.py --workload c5 --pulses $n --events $ev --steps 10 --warmup 3 --no-cpu-baseline --mode infer >> "$out" 2>> "${out%.jsonl}.err" || echo "{\"pulses\": $n, \"error\": \"rc=$?\"}" >> "$out"
done
python - "$out" <<'PY'
import json, sys
for l in open(sys.argv[1]):
    d = json.loads(l)
    if "error" in d: print(d); continue
    c = d["config"]
    nodes_per_s = c["nodes_per_batch"] / (d["ms_per_step"] * 1e-3)
    print(f'N={c["nodes_per_batch"]:8d}  events/s {d["value"]:12.0f}  Mnodes/s {nodes_per_s / 1e6:7.1f}  '
          f'ms/step {d["ms_per_step"]:8.3f}  e2e events/s {d["e2e"]["value"]:12.0f}  | {c["workload"][72:102]}')
PY
