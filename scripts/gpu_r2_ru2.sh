#!/bin/bash
mkdir -p gpurun_out
for dbg in 0 1 2 4 5 7; do
  echo "TB_RU_DEBUG=$dbg"; TB_RU_DEBUG=$dbg timeout 600 python scripts/rank_update_probe.py --prompts 512 2>&1 | sed -n 2p | cut -c1-700
done > gpurun_out/r02ru_debug.log 2>&1
cat gpurun_out/r02ru_debug.log
