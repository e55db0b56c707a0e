timeout 900 python -m pytest tests/test_gpu_sae.py tests/test_gpu_sweeps.py tests/test_gpu_targeting.py -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_enc2.json 2> gpurun_out/bench_enc2.err
