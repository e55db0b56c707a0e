set -x
timeout 900 python -m pytest tests/test_gpu_ll_fused.py tests/test_gpu_sweeps.py -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
python bench.py --steps 1 --warmup 1 --prompts 1024 --batch-prompts 1024 --workload random-control --seeds 2 > gpurun_out/bench_random.json 2> gpurun_out/bench_random.err
python bench.py --steps 1 --warmup 1 --prompts 1024 --batch-prompts 1024 --workload projection-vocab-shard > gpurun_out/bench_proj1.json 2> gpurun_out/bench_proj1.err
