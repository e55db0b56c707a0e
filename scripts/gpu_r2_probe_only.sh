#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_targeting.py tests/test_gpu_sae.py tests/test_gpu_ll_fused.py -q > gpurun_out/r02sel_pytest.log 2>&1; tail -4 gpurun_out/r02sel_pytest.log | cut -c1-300
timeout 600 python scripts/dense_ll_probe.py > gpurun_out/r02sel_dense_ll_pieces.jsonl 2> gpurun_out/r02sel_probe.err; echo "probe rc=$?"; grep -E "select|column sums|K3\)" gpurun_out/r02sel_dense_ll_pieces.jsonl | cut -c1-250; tail -3 gpurun_out/r02sel_probe.err
