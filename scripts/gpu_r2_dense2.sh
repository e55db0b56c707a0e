#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/dense_ll_probe.py > gpurun_out/r02x_dense_ll_pieces.jsonl 2> gpurun_out/r02x_probe.err; echo "probe rc=$?"; cat gpurun_out/r02x_dense_ll_pieces.jsonl
timeout 900 python -m pytest tests/test_gpu_ll_fused.py tests/test_gpu_lens.py tests/test_gpu_models.py -q > gpurun_out/r02x_pytest.log 2>&1; tail -4 gpurun_out/r02x_pytest.log
timeout 900 python bench.py --steps 2 --warmup 3 --no-workloads > gpurun_out/r02x_bench.json 2> gpurun_out/r02x_bench.err; echo "bench rc=$?"; tail -c 500 gpurun_out/r02x_bench.err
python - <<'PY'
import json
j = json.loads([l for l in open("gpurun_out/r02x_bench.json") if l.startswith("{")][-1])
print("value", j["value"], "e2e", j["e2e"]["value"], "ms", j["ms_per_step"])
print(j["roofline"]["achieved"], j["roofline"]["kernel_share_of_step"])
PY
