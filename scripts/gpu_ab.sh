set -x
mkdir -p gpurun_out
for rep in 1 2; do for c in 3 2; do
TT_SWEEP_CTAS=$c timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_b3_ctas${c}_$rep.log 2>&1
done; done
