#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/dense_ll_probe.py > gpurun_out/r02z2_dense_ll_pieces.jsonl 2> gpurun_out/r02z2_probe.err; echo "probe rc=$?"; cut -c1-250 gpurun_out/r02z2_dense_ll_pieces.jsonl; tail -3 gpurun_out/r02z2_probe.err
