set -x
timeout 900 python -m pytest tests/test_gpu_project.py tests/test_gpu_sweeps.py -m gpu -x -q > gpurun_out/pytest_proj.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_proj.log
python scripts/row_kernels_bench.py > gpurun_out/row_kernels.jsonl 2> gpurun_out/row_kernels.err
