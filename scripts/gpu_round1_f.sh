set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --ll-topk > gpurun_out/bench_ll.json 2> gpurun_out/bench_ll.err
