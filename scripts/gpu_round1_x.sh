set -x
timeout 900 python -m pytest tests/test_gpu_sae.py tests/test_gpu_sweeps.py tests/test_gpu_targeting.py tests/test_gpu_baselines.py -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
python scripts/k1b_probe.py > gpurun_out/k1b_probe.log 2>&1
