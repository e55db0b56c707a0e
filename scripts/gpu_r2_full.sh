#!/bin/bash
# whole -m gpu suite, default bench line (1 GPU), launch list + ncu --set full of the dominant lens kernel on a short run
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02r_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02r_pytest_gpu.log
tail -6 gpurun_out/r02r_pytest_gpu.log
python bench.py --steps 2 --warmup 3 > gpurun_out/r02r_bench.json 2> gpurun_out/r02r_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02r_bench.err
python - <<'PY'
import json
j = json.loads([l for l in open("gpurun_out/r02r_bench.json") if l.startswith("{")][-1])
print("value", j["value"], "e2e", j["e2e"]["value"], "ms", j["ms_per_step"])
print(j["roofline"]["hbm_path"])
for k, v in j.get("workloads", {}).items():
    print(k, {kk: v[kk] for kk in ("value", "ms_per_pass", "error") if kk in v})
PY
