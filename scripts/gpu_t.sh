set -x
timeout 900 python -m pytest $TB_TESTS -m gpu -x -q > gpurun_out/pytest_t.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_t.log
