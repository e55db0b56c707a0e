set -x
python -m pytest tests/test_gpu_ll_topk.py -m gpu -x -q > gpurun_out/pytest_ll.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ll.log
python scripts/lens_sweep.py --rows 75776 --seconds 3 > gpurun_out/lens_sweep.jsonl 2> gpurun_out/lens_sweep.err
python - <<'PY' > gpurun_out/occupancy.txt 2>&1
import torch
p = torch.cuda.get_device_properties(0)
print(p)
PY
