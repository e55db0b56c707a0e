#!/bin/bash
# rank-update lens: its tests, then the probe at the benchmark's shapes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rank_update.py -q > gpurun_out/r02ru_pytest.log 2>&1; tail -5 gpurun_out/r02ru_pytest.log | cut -c1-300
timeout 900 python scripts/rank_update_probe.py --prompts 512 > gpurun_out/r02ru_probe_last.jsonl 2> gpurun_out/r02ru_probe.err; echo "probe rc=$?"; head -3 gpurun_out/r02ru_probe_last.jsonl | cut -c1-700; tail -3 gpurun_out/r02ru_probe.err
