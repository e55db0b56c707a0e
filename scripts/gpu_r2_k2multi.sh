#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_project.py tests/test_gpu_sweeps.py -q > gpurun_out/r02k_pytest.log 2>&1; tail -12 gpurun_out/r02k_pytest.log | cut -c1-300
timeout 900 python bench.py --steps 1 --warmup 1 --prompts 512 --no-cpu-baseline --workloads projection_vocab_shard > gpurun_out/r02k_bench.json 2> gpurun_out/r02k_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02k_bench.err
python - <<'PY'
import json
j = json.loads([l for l in open("gpurun_out/r02k_bench.json") if l.startswith("{")][-1])
w = j["workloads"]["projection_vocab_shard"]
print({k: w[k] for k in ("value", "ms_per_pass", "k2_gbs", "phase_ms_this_rank") if k in w})
PY
