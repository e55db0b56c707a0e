set -x
timeout 900 python -m pytest tests/test_gpu_ll_fused.py -m gpu -x -q > gpurun_out/pytest_fused.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_fused.log
tail -1 gpurun_out/pytest_fused.log | grep -q "rc=0" && python bench.py --steps 1 --warmup 1 --no-cpu-baseline --ll-topk > gpurun_out/bench_ll.json 2> gpurun_out/bench_ll.err
