#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/k2_fused_check.py > gpurun_out/r02l_k2.jsonl 2> gpurun_out/r02l_k2.err; echo "k2 rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/r02l_k2.jsonl"):
    j = json.loads(l)
    if "kernel" in j: print(j["kernel"], j["rank"], round(j["ms"], 3), round(j["frac_of_measured_copy"], 3))
    else: print({k: (round(v, 6) if isinstance(v, float) else v) for k, v in j.items() if k.startswith(("n", "d", "r", "apply", "three"))})
PY
tail -5 gpurun_out/r02l_k2.err
timeout 600 python -m pytest tests/test_gpu_project.py tests/test_gpu_sweeps.py -q > gpurun_out/r02l_pytest.log 2>&1; tail -5 gpurun_out/r02l_pytest.log
