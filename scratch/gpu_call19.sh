set -x
DGIBBS_PROFILE=1 python bench.py --nbase 148 --nsamp 4 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c19_bench_small.log 2>&1
echo done
