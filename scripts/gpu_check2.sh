#!/bin/bash
# pregrasp bring-up: GPU parity tests (all), smoke, memcheck of a tiny pregrasp run
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest exit $?"; tail -60 gpurun_out/pytest_gpu2.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke2.log 2>&1; echo "smoke exit $?"; tail -5 gpurun_out/smoke2.log
