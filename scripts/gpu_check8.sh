#!/bin/bash
# f3 (TorchSDF compute_sdf on sm_100a): reference-kernel golden generation + comparison, SDF GPU tests
set -x
mkdir -p gpurun_out
timeout 600 python scripts/make_golden_sdf.py > gpurun_out/sdf_golden.log 2>&1; echo "golden exit $?"; tail -12 gpurun_out/sdf_golden.log
cp gpurun_out/sdf_reference_b200.npz tests/golden/ 2>/dev/null
timeout 900 python -m pytest tests/test_sdf_gpu.py tests/test_sdf_oracle.py -q -x > gpurun_out/pytest_sdf.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/pytest_sdf.log
