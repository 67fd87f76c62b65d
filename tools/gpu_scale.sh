#!/bin/bash
# One point of the scaling run: bench.py on N GPUs (torchrun, one rank per GPU); args: N workload tag
N=${1:-2}; W=${2:-c2}; TAG=${3:-rX}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
  python bench.py --workload $W --no-cpu-baseline > gpurun_out/${TAG}_scale_${W}_n${N}.json 2> gpurun_out/${TAG}_scale_${W}_n${N}.err
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py \
    --gpus $N --steps 10 --warmup 3 --workload $W > gpurun_out/${TAG}_scale_${W}_n${N}.json 2> gpurun_out/${TAG}_scale_${W}_n${N}.err
fi
echo "rc=$?"; tail -c 600 gpurun_out/${TAG}_scale_${W}_n${N}.json
