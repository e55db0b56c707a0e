#!/bin/bash
# rank-update lens: its tests, then the bench leg alone on a short headline
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rank_update.py -q > gpurun_out/r02ru_pytest.log 2>&1; tail -25 gpurun_out/r02ru_pytest.log | cut -c1-300
timeout 900 python bench.py --steps 1 --warmup 1 --prompts 512 --no-cpu-baseline --workloads rank_update > gpurun_out/r02ru_bench.json 2> gpurun_out/r02ru_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r02ru_bench.err
python - <<'PY'
import json
j = json.loads([l for l in open("gpurun_out/r02ru_bench.json") if l.startswith("{")][-1])
w = j["workloads"]["rank_update"]
print(json.dumps({k: v for k, v in w.items() if k not in ("drift_vs_dense_this_rank", "config", "not_the_reference_arithmetic")}, indent=0)[:3000])
PY
