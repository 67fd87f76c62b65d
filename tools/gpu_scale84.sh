#!/bin/bash
# 8-GPU box: bench at N = 8 and 4 (C4 headline + C2 block), no CPU legs
TAG=${1:-r2m}
mkdir -p gpurun_out
for N in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2967$N bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err; echo "bench n$N rc=$?"
done
python - <<P
import json
for N in (4,8):
    try:
        txt=open("gpurun_out/${TAG}_bench_n%d.json"%N).read()
        d=json.loads([l for l in txt.splitlines() if l.startswith("{")][-1])
    except Exception as e:
        print(N, "no json", e); continue
    for name, b in [("c4", d)] + list(d.get("also",{}).items()):
        print(N, name, "it/s %.1f ms %.3f build %.3f e2e %.1f"%(b["value"], b["ms_per_step"], b["schur_build_ms_per_iter"], b.get("e2e",{}).get("value",0)), {k: round(v,3) for k,v in b["phases_ms_per_iter"].items()})
P
