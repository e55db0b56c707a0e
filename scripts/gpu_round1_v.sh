set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -1 gpurun_out/pytest_gpu.log | grep -q "rc=0" || exit 1
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_split.json 2> gpurun_out/bench_split.err
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-ll-topk > gpurun_out/bench_split_noll.json 2> gpurun_out/bench_split_noll.err
