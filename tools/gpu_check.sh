#!/bin/bash
# One B200 box: GPU parity tests, smoke, bench lines, ncu launch lists + HBM-kernel counters.
# Everything lands in gpurun_out/ (merged back by gpurun); summaries are copied to profiles/ by hand.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "no-alloc alone rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 300 python bench.py --mode train --no-cpu-baseline > gpurun_out/bench_train_n1.json 2> gpurun_out/bench_train_n1.err; echo "train rc=$?"
timeout 300 python bench.py --workload c4 --events 256 --pulses 500 --no-cpu-baseline --mode infer > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "c4 rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 300 python tools/bench_knn.py > gpurun_out/bench_knn.jsonl 2> gpurun_out/bench_knn.err; echo "bench_knn rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --mode infer > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches_c4.csv \
  python bench.py --workload c4 --events 256 --pulses 500 --steps 2 --warmup 3 --no-cpu-baseline --mode infer > gpurun_out/ncu_launches_c4.log 2>&1; echo "ncu c4 launches rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 3000 --csv --log-file gpurun_out/train_launches.csv \
  python bench.py --mode train --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_train.log 2>&1; echo "ncu train rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none \
  -k regex:"segment_pool|global_variables|knn_kernel|coo_to_csr|linear_smallk" -c 60 --csv --log-file gpurun_out/hbm_c2.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --mode infer > gpurun_out/ncu_hbm_c2.log 2>&1; echo "ncu hbm c2 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none \
  -k regex:"segment_pool|global_variables|knn_kernel|coo_to_csr|linear_smallk" -c 60 --csv --log-file gpurun_out/hbm_c4.csv \
  python bench.py --workload c4 --events 256 --pulses 500 --steps 2 --warmup 3 --no-cpu-baseline --mode infer > gpurun_out/ncu_hbm_c4.log 2>&1; echo "ncu hbm c4 rc=$?"
tail -3 gpurun_out/pytest_gpu.log
cut -c1-300 gpurun_out/bench_n1.json
