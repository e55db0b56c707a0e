# the other two workloads of BASELINE.json on one GPU, reduced sizes
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload projection-vocab-shard --prompts 1024 > gpurun_out/bench_proj1.json 2> gpurun_out/bench_proj1.err
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --workload random-control --seeds 2 --prompts 2048 > gpurun_out/bench_random.json 2> gpurun_out/bench_random.err
