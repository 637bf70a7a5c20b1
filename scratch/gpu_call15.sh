set -x
timeout 900 python -m pytest tests/test_gpu_freq.py -m gpu -x -q > gpurun_out/c17_pytest_freq.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c17_pytest_freq.log
DGIBBS_PROFILE=1 python bench.py --nbase 148 --nsamp 4 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c17_bench_small.log 2>&1
echo done
