#!/bin/bash
# 2 GPUs: the NCCL parity test (log kept under profiles/), then the default bench line with its workload legs at N = 2
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_dist_nccl.py -m gpu -v > gpurun_out/r02zw_pytest_nccl_2gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r02zw_pytest_nccl_2gpu.log
tail -5 gpurun_out/r02zw_pytest_nccl_2gpu.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/r02zw_bench_2gpu.json 2> gpurun_out/r02zw_bench_2gpu.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/r02zw_bench_2gpu.err
python - <<'PY'
import json
try:
    j = json.loads([l for l in open("gpurun_out/r02zw_bench_2gpu.json") if l.startswith("{")][-1])
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
