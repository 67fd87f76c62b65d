#!/bin/bash
# ncu evidence for profiles/: launch list of the bench command, then one --set full capture of the two build kernels and the Cholesky.
# Run only after `python bench.py --steps 2 --warmup 1 --no-cpu-baseline` has exited 0 without ncu.
TAG=${1:-rX}
mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_short_${TAG}.json 2> gpurun_out/bench_short_${TAG}.err || { echo "bench failed"; tail -5 gpurun_out/bench_short_${TAG}.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${TAG}_launches_c2.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list_${TAG}.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_reduce_blocks|k_linearize|k_chol_dataflow" -c 5 -o gpurun_out/${TAG}_full_build_c2 -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full_${TAG}.log 2>&1
echo "full capture rc=$?"
ncu -i gpurun_out/${TAG}_full_build_c2.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_build_c2_raw.csv 2>/dev/null
ls -la gpurun_out | tail -12
