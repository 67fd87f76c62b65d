#!/bin/bash
# ncu --set full capture of the two build kernels (run only after the same command exited 0 without ncu)
TAG=${1:-r2a}; WL=${2:-c2}
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_reduce_blocks|k_linearize_packed" -c 4 -o gpurun_out/${TAG}_full_build_${WL} -f \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --workload ${WL} > gpurun_out/ncu_full_${TAG}.log 2>&1
echo "full capture rc=$?"
ncu -i gpurun_out/${TAG}_full_build_${WL}.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_build_${WL}_raw.csv 2>/dev/null
ls -la gpurun_out | tail -5
