set -x
timeout 900 python -m pytest tests/test_gpu_baselines.py -m gpu -x -q > gpurun_out/pytest_base.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_base.log
