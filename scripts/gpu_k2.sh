# row kernels: tests, the bandwidth table, per-kernel times of the tensor-core projection path at rank 64
timeout 600 python -m pytest tests/test_gpu_project.py tests/test_gpu_sweeps.py tests/test_gpu_lens.py tests/test_gpu_sae.py -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
timeout 600 python scripts/row_kernels_bench.py 262144 > gpurun_out/row_kernels_k2.jsonl 2> gpurun_out/row_kernels_k2.err
python scripts/k2_probe.py 262144 64 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/k2_launches.csv python scripts/k2_probe.py 262144 64 > gpurun_out/k2_ncu.log 2>&1
