#!/bin/bash
# full GPU suite (stop at first failure) + bench line without the CPU legs + e2e wall-clock split
TAG=${1:-f}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.log
tail -4 gpurun_out/${TAG}_pytest.log
timeout 600 python bench.py --no-cpu-baseline --no-e2e-cpp > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
python - <<P
import json
d=json.load(open("gpurun_out/${TAG}_bench.json"))
def show(n,d):
    print(n, "it/s %.1f ms %.3f build %.3f e2e %.1f"%(d["value"], d["ms_per_step"], d["schur_build_ms_per_iter"], d.get("e2e",{}).get("value",0)))
    print("   phases", {k: round(v,3) for k,v in d["phases_ms_per_iter"].items()})
show("head", d)
for k,v in d.get("also",{}).items(): show(k, v)
P
BAGPU_TIMING=1 python tools/diag_e2e.py c4 10 pinned 2>&1 | tail -28
