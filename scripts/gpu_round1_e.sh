set -x
for m in 4 2; do
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --cluster-mode $m > gpurun_out/bench_mode$m.json 2> gpurun_out/bench_mode$m.err
done
