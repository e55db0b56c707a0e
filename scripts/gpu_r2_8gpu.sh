#!/bin/bash
# 8 GPUs: the default bench line with its workload legs at N = 8
mkdir -p gpurun_out
nvidia-smi -L
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 2 --warmup 3 > gpurun_out/r02zy_bench_8gpu.json 2> gpurun_out/r02zy_bench_8gpu.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/r02zy_bench_8gpu.err
python - <<'PY'
import json
try:
    j = json.loads([l for l in open("gpurun_out/r02zy_bench_8gpu.json") if l.startswith("{")][-1])
    print("value", j["value"], "e2e", j["e2e"]["value"])
    for k, v in j.get("workloads", {}).items():
        print(k, {kk: v[kk] for kk in ("value", "ms_per_pass", "error", "wall_s") if kk in v})
        if "alternative_k2_row_sharded" in v:
            print("   alt", v["alternative_k2_row_sharded"]["value"], v["alternative_k2_row_sharded"]["phase_ms_this_rank"])
        if "phase_ms_this_rank" in v:
            print("   ", v["phase_ms_this_rank"], v.get("lens_tflops_per_gpu"))
except Exception as e:
    print("parse failed", e)
PY
