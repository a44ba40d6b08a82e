#!/bin/bash
# Round-end artefacts for profiles/: launch lists (C2, C4), ncu --set full captures of the edge
# kernel and the tcgen05 GEMM, the training line, the kernel timelines (profiling build).
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_c4.csv \
  python bench.py --workload c4 --events 256 --pulses 500 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches_c4.log 2>&1; echo "ncu c4 launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:edge_mlp_aggregate_pair -s 5 -c 3 \
  -o gpurun_out/r02_edge_pair -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_edge_pair.log 2>&1; echo "ncu edge rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_linear_kernel -s 15 -c 7 \
  -o gpurun_out/r02_tc_linear -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_tc_linear.log 2>&1; echo "ncu linear rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none \
  -k regex:"segment_pool|global_variables|knn_kernel|coo_to_csr|pool_finalize" -c 60 --csv --log-file gpurun_out/hbm_c2.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-riders > gpurun_out/ncu_hbm_c2.log 2>&1; echo "ncu hbm rc=$?"
timeout 300 python bench.py --mode train --no-cpu-baseline > gpurun_out/bench_train_n1.json 2> gpurun_out/bench_train_n1.err; echo "train rc=$?"
timeout 300 python bench.py --workload c4 --events 256 --pulses 500 --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "c4 rc=$?"
GN_B200_DEBUG_LIB=1 GN_EDGE_DEBUG=1 timeout 120 python tools/trace_edge.py c2 1 24 > gpurun_out/trace_edge_c2_l1.txt 2>/dev/null; echo "trace edge rc=$?"
GN_B200_DEBUG_LIB=1 GN_EDGE_DEBUG=1 timeout 120 python tools/time_edge.py c2 > gpurun_out/edge_counters_c2.jsonl 2>/dev/null
GN_B200_DEBUG_LIB=1 GN_EDGE_DEBUG=1 timeout 120 python tools/time_edge.py c4 > gpurun_out/edge_counters_c4.jsonl 2>/dev/null
GN_B200_DEBUG_LIB=1 GN_TC_DEBUG=1 timeout 120 python tools/trace_linear.py > gpurun_out/trace_linear_node_gemm.txt 2>/dev/null; echo "trace linear rc=$?"
timeout 60 ./build/tmem_ld_bench > gpurun_out/tmem_ld_bench.txt 2>&1; timeout 60 ./build/mma2_bench > gpurun_out/mma2_bench.txt 2>&1
cut -c1-200 gpurun_out/bench_train_n1.json; echo; cut -c1-200 gpurun_out/bench_c4.json
