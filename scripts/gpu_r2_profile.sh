#!/bin/bash
# round 2 evidence: tests of the new dense-probs mode, then ncu: launch list of a short bench (kernel shares), --set full
# of the dominant lens kernel, of the encode kernel in its cluster mode, and of the one-pass K2 kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_lens.py tests/test_gpu_models.py tests/test_gpu_baselines.py -q > gpurun_out/r02j_pytest.log 2>&1; tail -4 gpurun_out/r02j_pytest.log
SMALL="bench.py --steps 1 --warmup 1 --prompts 296 --batch-prompts 296 --no-cpu-baseline --no-workloads"
python $SMALL > gpurun_out/plain_small.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_latest.csv python $SMALL > gpurun_out/ncu_launch.log 2>&1
python $SMALL > gpurun_out/plain_small2.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:LensLLEpi -s 2 -c 2 -o gpurun_out/lensll_latest python $SMALL > gpurun_out/ncu_full.log 2>&1
python scripts/encode_probe.py --S 16384 --reps 1 > gpurun_out/plain16.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tb_umma_gemm_kernel -s 1 -c 1 -o gpurun_out/r02j_encode16k_clusters python scripts/encode_probe.py --S 16384 --reps 1 > gpurun_out/ncu16.log 2>&1
python scripts/k2_fused_ncu.py > gpurun_out/plaink2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:tb_lowrank_cluster_kernel -s 1 -c 1 -o gpurun_out/r02j_k2_one_pass python scripts/k2_fused_ncu.py > gpurun_out/ncuk2.log 2>&1
ls -la gpurun_out | grep -E "ncu-rep|launches_latest"
