#!/bin/bash
# final visit of round 2: full GPU suite, smoke, the default bench line (all legs), then (after those exited 0 without a
# profiler) the launch lists and the ncu --set full pages of the build kernels on C2 and C4
TAG=${1:-r2q}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
for WL in c4 c2; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${TAG}_launches_${WL}.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --also "" --workload $WL > gpurun_out/${TAG}_ncu_list_${WL}.log 2>&1; echo "launch list $WL rc=$?"
  python tools/launch_summary.py gpurun_out/${TAG}_launches_${WL}.csv > gpurun_out/${TAG}_launch_summary_${WL}.txt 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_reduce_blocks|k_linearize_packed" -c 4 -o gpurun_out/${TAG}_full_build_${WL} -f \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --also "" --workload $WL > gpurun_out/${TAG}_ncu_full_${WL}.log 2>&1; echo "full $WL rc=$?"
  ncu -i gpurun_out/${TAG}_full_build_${WL}.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_build_${WL}_raw.csv 2>/dev/null
  rm -f gpurun_out/${TAG}_full_build_${WL}.ncu-rep
done
python - <<P
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
def show(n,d):
    print(n, "it/s %.1f ms %.3f build %.3f e2e %.1f"%(d["value"], d["ms_per_step"], d["schur_build_ms_per_iter"], d.get("e2e",{}).get("value",0)), "e2e_cpp", d.get("e2e_cpp"))
    print("   phases", {k: round(v,3) for k,v in d["phases_ms_per_iter"].items()})
    r=d.get("roofline");
    if r: print("   roofline", r["kernel"], round(r["frac"],3), "build frac", round(r["build"]["frac"],3))
    print("   cpu", d.get("cpu_baseline",{}).get("value"), "parity", {k: d.get("parity",{}).get(k) for k in ("max_rel_diff_costs","max_rel_diff_steps")})
show("head", d)
for k,v in d.get("also",{}).items(): show(k, v)
P
