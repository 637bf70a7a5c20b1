set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/c43_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c43_pytest.log
DGIBBS_PROFILE=1 timeout 600 python bench.py --nbase 296 --nsamp 10 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/c43_bench_small.log 2>&1; echo "rc=$?" >> gpurun_out/c43_bench_small.log
