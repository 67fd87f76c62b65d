#!/bin/bash
# bench.py on every workload (N = 1) plus the reference arm; arg: tag
TAG=${1:-rX}
mkdir -p gpurun_out
timeout 300 python bench.py > gpurun_out/${TAG}_bench_c2.json 2> gpurun_out/${TAG}_bench_c2.err; echo "c2 rc=$?"
for w in c1 c3 c4 c5; do timeout 300 python bench.py --workload $w --no-cpu-baseline > gpurun_out/${TAG}_bench_$w.json 2> gpurun_out/${TAG}_bench_$w.err; echo "$w rc=$?"; done
timeout 300 python bench.py --impl reference --steps 2 --warmup 0 > gpurun_out/${TAG}_bench_reference_c2.json 2> gpurun_out/${TAG}_bench_reference_c2.err; echo "ref rc=$?"
