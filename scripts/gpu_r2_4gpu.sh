#!/bin/bash
# 4 GPUs: the default bench line with its workload legs at N = 4
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 4 --steps 2 --warmup 3 > gpurun_out/r02zv_bench_4gpu.json 2> gpurun_out/r02zv_bench_4gpu.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    j = json.loads([l for l in open("gpurun_out/r02zv_bench_4gpu.json") if l.startswith("{")][-1])
    print("value", j["value"], "e2e", j["e2e"]["value"])
    for k, v in j.get("workloads", {}).items():
        print(k, {kk: v[kk] for kk in ("value", "ms_per_pass", "error") if kk in v})
except Exception as e:
    print("parse failed", e)
PY
