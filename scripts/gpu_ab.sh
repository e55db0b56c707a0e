# A/B of two library builds in one session, then the tests that exercise the epilogues' shared memory
python scripts/ab_lib.py scripts/probes/libtb_old.so taboo_brittleness_b200/libtb_b200.so --rounds 2 > gpurun_out/ab_lib.jsonl 2> gpurun_out/ab_lib.err
timeout 900 python -m pytest tests/test_gpu_lens.py tests/test_gpu_ll_fused.py tests/test_gpu_ll_topk.py tests/test_gpu_project.py tests/test_gpu_gemm.py -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
