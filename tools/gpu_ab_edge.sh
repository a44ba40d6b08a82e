#!/bin/bash
for l in lib libDGN_E4_NOCLAMP libDGN_E4_KPLD libDGN_E4_NOCLAMPDGN_E4_KPLD lib; do
  echo "== $l"; GN_B200_LIB=$PWD/build/$l.so timeout 200 python tools/time_edge.py c2 2>/dev/null | head -1 | cut -c1-200
done
for k in 0 1; do
  echo "== kpb1=$k"
  if [ $k = 1 ]; then export GN_TC_KPB1=1; else unset GN_TC_KPB1; fi
  timeout 300 python tools/insitu_kernels.py 2>/dev/null | grep "tc_linear\|sum of"
done
