#!/bin/bash
# one --set full capture of a kernel (regex) in the build-only driver; args: regex workload tag [skip]
RE=$1; W=${2:-c2}; TAG=${3:-x}; SKIP=${4:-1}
mkdir -p gpurun_out
python tools/diag_build.py $W 3 > gpurun_out/build_${TAG}.log 2>&1 || { tail -5 gpurun_out/build_${TAG}.log; exit 1; }
cat gpurun_out/build_${TAG}.log | tail -1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RE" -s $SKIP -c 1 -o gpurun_out/${TAG} -f python tools/diag_build.py $W 3 > gpurun_out/ncu_${TAG}.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/${TAG}.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
ncu -i gpurun_out/${TAG}.ncu-rep --page details 2>/dev/null | grep -E "Duration|DRAM Throughput|L1/TEX Hit|L2 Hit|Registers Per|Achieved Occupancy|Theoretical Occ|Issue Slots Busy|Executed Ipc|Mem Busy|Max Bandwidth|L2 Cache Throughput|L1/TEX Cache Throughput|No Eligible|Stall" | head -30
