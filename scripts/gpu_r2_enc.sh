#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sae.py tests/test_gpu_fullsize.py -q > gpurun_out/r02q_pytest.log 2>&1; tail -6 gpurun_out/r02q_pytest.log
for sub in 1 2 4; do timeout 300 python scripts/encode_probe.py --S 16384 --sub $sub 2>&1 | tail -1; done | tee gpurun_out/r02q_encode_16k.jsonl
for sub in 1 2 4; do timeout 300 python scripts/encode_probe.py --S 131072 --rows 37888 --sub $sub 2>&1 | tail -1; done | tee gpurun_out/r02q_encode_131k.jsonl
python scripts/encode_probe.py --S 16384 --sub 2 --reps 1 > gpurun_out/plain16.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tb_umma_gemm_kernel -s 1 -c 1 -o gpurun_out/r02q_encode16k_chunks python scripts/encode_probe.py --S 16384 --sub 2 --reps 1 > gpurun_out/ncu16.log 2>&1
