#!/bin/bash
mkdir -p gpurun_out
export GN_B200_DEBUG_LIB=1 GN_NO_FLUSH=1 GN_TC_DEBUG=1
for mc in 4 1; do for h in 0 24; do
echo "=== MAX_CLUSTER=$mc HACK=$h"
GN_TC_MAX_CLUSTER=$mc GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 704 256 2>&1 | grep -v "^item" | head -34
done; done > gpurun_out/lin_stages.txt 2>&1
cat gpurun_out/lin_stages.txt
