#!/bin/bash
# launch list (gpu__time_duration per launch) of a short run; args: workload iters scale tag
W=${1:-c2}; IT=${2:-2}; SC=${3:-1.0}; TAG=${4:-x}
python tools/diag.py $W $IT $SC > gpurun_out/diag_${TAG}.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_${TAG}.csv python tools/diag.py $W $IT $SC > gpurun_out/ncu_${TAG}.log 2>&1
tail -3 gpurun_out/diag_${TAG}.log
