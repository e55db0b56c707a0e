set -x
timeout 600 python -m pytest tests/test_gpu_lens.py tests/test_gpu_ll_topk.py tests/test_gpu_gemm.py -m gpu -x -q > gpurun_out/pytest_mc.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_mc.log
tail -3 gpurun_out/pytest_mc.log | grep -q "rc=0" && timeout 600 python scripts/lens_sweep.py --rows 75776 --seconds 2.5 --chunks 2,4,8,10,20 --no-cublas > gpurun_out/lens_sweep_mc.jsonl 2> gpurun_out/lens_sweep_mc.err
