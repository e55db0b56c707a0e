set -x
timeout 900 python -m pytest tests/test_gpu_ll_fused.py tests/test_gpu_ll_topk.py tests/test_gpu_sweeps.py -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
tail -1 gpurun_out/pytest_t.log | grep -q "rc=0" || exit 1
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_ov.json 2> gpurun_out/bench_ov.err
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-overlap > gpurun_out/bench_noov.json 2> gpurun_out/bench_noov.err
