#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/k2_profile.py --ranks 16,64 > gpurun_out/r02d_k2_profile.jsonl 2> gpurun_out/r02d_k2_profile.err; echo "prof rc=$?"
cat gpurun_out/r02d_k2_profile.jsonl; tail -3 gpurun_out/r02d_k2_profile.err
timeout 300 python scripts/k2_fused_check.py --skip-check > gpurun_out/r02d_k2_fused.jsonl 2>&1; cat gpurun_out/r02d_k2_fused.jsonl
timeout 600 python -m pytest tests/test_gpu_configs.py tests/test_gpu_project.py -q > gpurun_out/r02d_pytest.log 2>&1; tail -8 gpurun_out/r02d_pytest.log
