# wait profile of the lens kernels (scripts/wait_profile.py), a subset of the GPU tests, the default bench
python scripts/wait_profile.py --rows 75776 --modes 5,2,1 > gpurun_out/wait_prof_elect.jsonl 2> gpurun_out/wait_prof_elect.err
timeout 900 python -m pytest tests/test_gpu_lens.py tests/test_gpu_ll_fused.py tests/test_gpu_ll_topk.py tests/test_gpu_gemm.py tests/test_gpu_sae.py -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_elect.json 2> gpurun_out/bench_elect.err
