#!/bin/bash
# round 2, second GPU call: the whole -m gpu suite (no -x), then the ncu --set full captures of K1 at both widths
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02b_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02b_pytest_gpu.log
tail -15 gpurun_out/r02b_pytest_gpu.log
python scripts/encode_probe.py --S 16384 --reps 1 > gpurun_out/plain16.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tb_umma_gemm_kernel -s 1 -c 1 -o gpurun_out/r02b_encode16k python scripts/encode_probe.py --S 16384 --reps 1 > gpurun_out/ncu16.log 2>&1
python scripts/encode_probe.py --S 131072 --rows 37888 --reps 1 > gpurun_out/plain131.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tb_umma_gemm_kernel -s 1 -c 1 -o gpurun_out/r02b_encode131k python scripts/encode_probe.py --S 131072 --rows 37888 --reps 1 > gpurun_out/ncu131.log 2>&1
tail -3 gpurun_out/ncu16.log gpurun_out/ncu131.log
ls -la gpurun_out | tail -12
