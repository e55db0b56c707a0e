#!/bin/bash
# final state of round 2: whole -m gpu suite, smoke(), default bench line (1 GPU, with the workload legs)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r02zv_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02zv_pytest_gpu.log
tail -6 gpurun_out/r02zv_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02zv_smoke.log 2>&1; tail -2 gpurun_out/r02zv_smoke.log
python bench.py > gpurun_out/r02zv_bench.json 2> gpurun_out/r02zv_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02zv_bench.err
python - <<'PY'
import json
j = json.loads([l for l in open("gpurun_out/r02zv_bench.json") if l.startswith("{")][-1])
print("value", j["value"], "e2e", j["e2e"]["value"], "ms", j["ms_per_step"], "steps", j["steps"], "warmup", j["warmup"])
print(j["roofline"]["hbm_path"]); print(j["clocks"]); print(j["cpu_baseline"])
for k, v in j.get("workloads", {}).items():
    print(k, {kk: v[kk] for kk in ("value", "ms_per_pass", "error", "speedup_over_dense") if kk in v})
PY
