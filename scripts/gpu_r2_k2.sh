#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/k2_fused_check.py > gpurun_out/r02c_k2_fused.jsonl 2> gpurun_out/r02c_k2_fused.err; echo "k2 rc=$?"
cat gpurun_out/r02c_k2_fused.jsonl; tail -5 gpurun_out/r02c_k2_fused.err
timeout 300 python scripts/encode_probe.py --S 16384 > gpurun_out/r02c_encode_16k.json 2>&1; cat gpurun_out/r02c_encode_16k.json | tail -3
timeout 300 python scripts/encode_probe.py --S 131072 --rows 37888 > gpurun_out/r02c_encode_131k.json 2>&1; tail -2 gpurun_out/r02c_encode_131k.json
timeout 900 python -m pytest tests/test_gpu_sae.py tests/test_gpu_configs.py tests/test_gpu_project.py tests/test_gpu_lens.py -q > gpurun_out/r02c_pytest.log 2>&1; tail -8 gpurun_out/r02c_pytest.log
