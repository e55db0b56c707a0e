#!/bin/bash
mkdir -p gpurun_out
timeout 900 python scripts/rank_update_probe.py --prompts 512 > gpurun_out/r02ru_probe_desc.jsonl 2> gpurun_out/r02ru_probe.err; echo "probe rc=$?"; head -3 gpurun_out/r02ru_probe_desc.jsonl | cut -c1-700; tail -3 gpurun_out/r02ru_probe.err
