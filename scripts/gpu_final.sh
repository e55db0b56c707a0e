# round-end sequence: all GPU tests, default bench, row-kernel table, smoke
set -x
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err
timeout 600 python scripts/row_kernels_bench.py 262144 > gpurun_out/row_kernels_final.jsonl 2> gpurun_out/row_kernels_final.err
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
