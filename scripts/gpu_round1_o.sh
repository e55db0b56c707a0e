set -x
timeout 900 python -m pytest tests/test_gpu_sweeps.py tests/test_gpu_sae.py tests/test_gpu_baselines.py -m gpu -x -q > gpurun_out/pytest_sweeps.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_sweeps.log
