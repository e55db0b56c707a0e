#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/rank_update_probe.py --prompts 512 > gpurun_out/r02ru_probe_final.jsonl 2> gpurun_out/r02ru_probe_final.err; echo "probe rc=$?"; head -3 gpurun_out/r02ru_probe_final.jsonl | cut -c1-600; tail -3 gpurun_out/r02ru_probe_final.jsonl; tail -3 gpurun_out/r02ru_probe_final.err
python scripts/rank_update_ncu.py > gpurun_out/plain_ru.log 2>&1 && cat gpurun_out/plain_ru.log &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:rank_update_lens -s 2 -c 2 -o gpurun_out/r02ru_update_kernel python scripts/rank_update_ncu.py > gpurun_out/ncu_ru.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_ru.log
