#!/bin/bash
mkdir -p gpurun_out
export GN_B200_DEBUG_LIB=1
for h in 0 8 16 24; do
    echo "== HACK=$h cold / warm (lda 1088)"
    GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 704 256 2>&1 | head -1
    GN_NO_FLUSH=1 GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 704 256 2>&1 | head -1
    GN_NO_FLUSH=1 GN_TC_HACK=$h timeout 120 python tools/trace_linear.py 61443 336 1088 2>&1 | head -1
done 2>&1 | tee gpurun_out/lin_ablate4.txt
GN_NO_FLUSH=1 GN_TC_DEBUG=1 timeout 120 python tools/trace_linear.py 61443 704 256 > gpurun_out/lin_trace3.txt 2>&1; head -8 gpurun_out/lin_trace3.txt
