#!/bin/bash
# round 2, first GPU call: the whole -m gpu suite, the default bench line (headline + workload legs), K1 alone at
# both SAE widths and its ncu --set full captures
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02a_pytest_gpu.log
tail -5 gpurun_out/r02a_pytest_gpu.log
python bench.py --steps 2 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r02a_bench.err
python scripts/encode_probe.py --S 16384 > gpurun_out/r02a_encode_16k.json 2>&1
python scripts/encode_probe.py --S 131072 --rows 37888 > gpurun_out/r02a_encode_131k.json 2>&1
cat gpurun_out/r02a_encode_16k.json gpurun_out/r02a_encode_131k.json
python scripts/encode_probe.py --S 16384 --reps 1 > gpurun_out/plain16.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tb_umma_gemm_kernel -s 1 -c 1 -o gpurun_out/r02a_encode16k python scripts/encode_probe.py --S 16384 --reps 1 > gpurun_out/ncu16.log 2>&1
python scripts/encode_probe.py --S 131072 --rows 37888 --reps 1 > gpurun_out/plain131.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tb_umma_gemm_kernel -s 1 -c 1 -o gpurun_out/r02a_encode131k python scripts/encode_probe.py --S 131072 --rows 37888 --reps 1 > gpurun_out/ncu131.log 2>&1
ls -la gpurun_out | tail -20
