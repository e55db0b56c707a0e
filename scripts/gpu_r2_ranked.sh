#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sae.py tests/test_gpu_sweeps.py tests/test_gpu_configs.py tests/test_gpu_ll_fused.py tests/test_gpu_targeting.py -q > gpurun_out/r02h_pytest.log 2>&1; tail -8 gpurun_out/r02h_pytest.log
timeout 600 python scripts/row_kernels_bench.py 262144 > gpurun_out/r02h_row_kernels_hbm.jsonl 2> gpurun_out/r02h_row.err; tail -16 gpurun_out/r02h_row_kernels_hbm.jsonl; tail -3 gpurun_out/r02h_row.err
