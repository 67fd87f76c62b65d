#!/bin/bash
# round-2 visit A: parity tests, bench C2 / C4 (no CPU leg), launch lists, ncu --set full of the build kernels
TAG=${1:-r2a}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -8 gpurun_out/${TAG}_pytest.log
for WL in c2 c4; do
  timeout 300 python bench.py --workload $WL --no-cpu-baseline > gpurun_out/${TAG}_bench_${WL}.json 2> gpurun_out/${TAG}_bench_${WL}.err; echo "bench $WL rc=$?"
done
for WL in c1 c3 c5; do
  timeout 300 python bench.py --workload $WL --no-cpu-baseline > gpurun_out/${TAG}_bench_${WL}.json 2> gpurun_out/${TAG}_bench_${WL}.err; echo "bench $WL rc=$?"
done
for WL in c2 c4; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${TAG}_launches_${WL}.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --workload $WL > gpurun_out/${TAG}_ncu_list_${WL}.log 2>&1; echo "launch list $WL rc=$?"
  python tools/launch_summary.py gpurun_out/${TAG}_launches_${WL}.csv > gpurun_out/${TAG}_launch_summary_${WL}.txt 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_reduce_blocks|k_linearize_packed|k_residual" -c 6 -o gpurun_out/${TAG}_full_build_${WL} -f \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --workload $WL > gpurun_out/${TAG}_ncu_full_${WL}.log 2>&1; echo "full $WL rc=$?"
  ncu -i gpurun_out/${TAG}_full_build_${WL}.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_build_${WL}_raw.csv 2>/dev/null
done
python - <<'P'
import json,sys
tag=sys.argv[1] if len(sys.argv)>1 else "r2a"
for n in ("c2","c4","c1","c3","c5"):
    try:
        d=json.load(open("gpurun_out/r2a_bench_%s.json"%n))
        print(n, d["value"], d["ms_per_step"], d.get("schur_build_ms_per_iter"), d.get("phases_ms_per_iter"), d.get("final_cost"), d.get("e2e",{}).get("value"), d.get("roofline"))
    except Exception as e: print(n, "no json", e)
P
ls -la gpurun_out | tail -30
