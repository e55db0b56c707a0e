#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sae.py tests/test_gpu_sweeps.py tests/test_gpu_ll_fused.py tests/test_gpu_fullsize.py tests/test_gpu_rank_update.py -q > gpurun_out/r02m_pytest.log 2>&1; tail -5 gpurun_out/r02m_pytest.log | cut -c1-300
timeout 900 python bench.py --steps 2 --warmup 3 --no-workloads --no-cpu-baseline > gpurun_out/r02m_bench.json 2> gpurun_out/r02m_bench.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r02m_bench.err
python - <<'PY'
import json
j = json.loads([l for l in open("gpurun_out/r02m_bench.json") if l.startswith("{")][-1])
print("value", j["value"], "e2e", j["e2e"]["value"], "ms", j["ms_per_step"])
print(j["roofline"]["hbm_path"])
PY
