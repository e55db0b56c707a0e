#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02k_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02k_pytest_gpu.log
tail -6 gpurun_out/r02k_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02k_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r02k_smoke.log
