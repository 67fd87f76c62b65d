#!/bin/bash
# backward-chain check: solver trace on C2 / C4, then the parity tests and the bench line
TAG=${1:-ch}
mkdir -p gpurun_out
BAGPU_CHOL_TRACE=gpurun_out/${TAG}_chol_c2.bin timeout 120 python tools/diag.py c2 2 1.0 > gpurun_out/${TAG}_diag_c2.log 2>&1; echo "diag c2 rc=$?"
python tools/chol_trace.py gpurun_out/${TAG}_chol_c2.bin | head -9
BAGPU_CHOL_TRACE=gpurun_out/${TAG}_chol_c4.bin timeout 120 python tools/diag.py c4 2 1.0 > gpurun_out/${TAG}_diag_c4.log 2>&1; echo "diag c4 rc=$?"
python tools/chol_trace.py gpurun_out/${TAG}_chol_c4.bin | head -9

bash tools/gpu_quick.sh ${TAG}
