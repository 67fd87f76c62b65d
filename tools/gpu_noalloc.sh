#!/bin/bash
# the allocation-free entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
TAG=${1:-na}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/${TAG}_alone.log 2>&1; echo "alone rc=$?"
tail -15 gpurun_out/${TAG}_alone.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "suite rc=$?"
tail -5 gpurun_out/${TAG}_pytest.log
