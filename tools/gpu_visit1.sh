#!/bin/bash
# GPU visit: parity tests (stop at the first failure), then bench lines without the CPU leg.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -30 gpurun_out/pytest.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench c2 rc=$?"
timeout 300 python bench.py --workload c4 --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "bench c4 rc=$?"
tail -3 gpurun_out/bench_c2.err gpurun_out/bench_c4.err
python - <<'P'
import json
for n in ("c2","c4"):
    try:
        d=json.load(open("gpurun_out/bench_%s.json"%n))
        print(n, d["value"], d["ms_per_step"], d["schur_build_ms_per_iter"], d["phases_ms_per_iter"], d["final_cost"], d["e2e"]["value"])
    except Exception as e: print(n, "no json", e)
P
