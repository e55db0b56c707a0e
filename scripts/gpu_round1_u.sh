set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -1 gpurun_out/pytest_gpu.log | grep -q "rc=0" || exit 1
timeout 600 python scripts/lens_sweep.py --rows 75776 --seconds 2.5 --chunks 2,4 --modes 2 --no-cublas > gpurun_out/lens_sweep_log2.jsonl 2> gpurun_out/lens_sweep_log2.err
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_log2.json 2> gpurun_out/bench_log2.err
