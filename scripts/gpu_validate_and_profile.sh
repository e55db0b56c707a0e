set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
python bench.py --steps 2 --warmup 3 --no-ll-topk --no-cpu-baseline > gpurun_out/bench_noll.json 2> gpurun_out/bench_noll.err
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > gpurun_out/clocks.csv &
SMI=$!
python bench.py --steps 1 --warmup 1 --prompts 296 --batch-prompts 296 --no-cpu-baseline > gpurun_out/plain_small.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_latest.csv python bench.py --steps 1 --warmup 1 --prompts 296 --batch-prompts 296 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
python bench.py --steps 1 --warmup 1 --prompts 296 --batch-prompts 296 --no-cpu-baseline > gpurun_out/plain_small2.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:LensLLEpi -s 2 -c 2 -o gpurun_out/lensll_latest python bench.py --steps 1 --warmup 1 --prompts 296 --batch-prompts 296 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
kill $SMI
