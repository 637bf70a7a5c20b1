set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c20_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c20_pytest.log
DGIBBS_PROFILE=1 python bench.py --nbase 148 --nsamp 4 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c20_bench_small.log 2>&1
timeout 900 python scratch/perf_extra.py > gpurun_out/c20_perf_extra.log 2>&1
timeout 300 python bench.py --config cfg3 --nbase 2048 --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/c20_bench_cfg3.log 2>&1
echo done
