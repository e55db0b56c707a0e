# K2 tensor-core path: tests, the row-kernel bandwidth table, per-kernel times at rank 64, one full capture of pass 2
timeout 600 python -m pytest tests/test_gpu_project.py tests/test_gpu_sweeps.py -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
timeout 600 python scripts/row_kernels_bench.py 262144 > gpurun_out/row_kernels_k2.jsonl 2> gpurun_out/row_kernels_k2.err
python scripts/k2_probe.py 262144 64 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/k2_launches.csv python scripts/k2_probe.py 262144 64 > gpurun_out/k2_ncu.log 2>&1
python scripts/k2_probe.py 262144 64 && ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:ProjApplyEpi -s 1 -c 1 -o gpurun_out/k2_apply python scripts/k2_probe.py 262144 64 > gpurun_out/k2_ncu_full.log 2>&1
