set -x
timeout 900 python -m pytest tests/test_gpu_lens.py tests/test_gpu_project.py tests/test_gpu_sae.py -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
python scripts/row_kernels_bench.py > gpurun_out/row_kernels.jsonl 2> gpurun_out/row_kernels.err
