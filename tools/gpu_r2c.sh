#!/bin/bash
# visit: parity tests, the default bench line (C4 headline + C2 block, CPU legs, parity block)
TAG=${1:-r2c}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -4 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/${TAG}_bench.err
python - <<P
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
def show(n,d):
    print(n, "it/s %.1f ms %.3f build %.3f e2e %.1f"%(d["value"], d["ms_per_step"], d["schur_build_ms_per_iter"], d.get("e2e",{}).get("value",0)))
    print("   phases", {k: round(v,3) for k,v in d["phases_ms_per_iter"].items()})
    r=d.get("roofline");
    if r: print("   roofline", r["kernel"], round(r["frac"],3), "build frac", round(r["build"]["frac"],3))
    print("   cpu", d.get("cpu_baseline"), "parity", d.get("parity"))
show("head", d)
for k,v in d.get("also",{}).items(): show(k, v)
P
