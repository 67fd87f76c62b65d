#!/bin/bash
TAG=${1:-r2e}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_multi_gpu.py -m gpu -q -k "${2:-two_gpu}" > gpurun_out/${TAG}_pytest_mg.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_mg.log
tail -40 gpurun_out/${TAG}_pytest_mg.log
