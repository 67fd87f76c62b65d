#!/bin/bash
# 2-GPU visit: the multi-GPU parity tests, then bench lines at N=2 (C4 headline + C2 block)
TAG=${1:-r2d}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${TAG}_smi.txt
timeout 1500 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/${TAG}_pytest_mg.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_mg.log
tail -30 gpurun_out/${TAG}_pytest_mg.log
BAGPU_TIMING=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_n2.json 2> gpurun_out/${TAG}_bench_n2.err; echo "bench n2 rc=$?"
grep -a "peer" gpurun_out/${TAG}_bench_n2.err | head -6
python - <<P
import json
d=json.loads([l for l in open("gpurun_out/${TAG}_bench_n2.json").read().splitlines() if l.startswith("{")][-1])
def show(n,d):
    print(n, "it/s %.1f ms %.3f build %.3f e2e %.1f"%(d["value"], d["ms_per_step"], d["schur_build_ms_per_iter"], d.get("e2e",{}).get("value",0)), "cost", d["final_cost"])
    print("   phases", {k: round(v,3) for k,v in d["phases_ms_per_iter"].items()})
show("head", d)
for k,v in d.get("also",{}).items(): show(k, v)
P
