set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -1 gpurun_out/pytest_gpu.log | grep -q "rc=0" || exit 1
timeout 600 python scripts/lens_sweep.py --rows 75776 --seconds 2.5 --chunks 1,2,3,4,5,6 --modes 2 --no-cublas > gpurun_out/lens_sweep_m2.jsonl 2> gpurun_out/lens_sweep_m2.err
for m in 2 4; do
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --cluster-mode $m > gpurun_out/bench_ls_mode$m.json 2> gpurun_out/bench_ls_mode$m.err
done
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --cluster-mode 2 --no-ll-topk > gpurun_out/bench_ls_mode2_noll.json 2> gpurun_out/bench_ls_mode2_noll.err
