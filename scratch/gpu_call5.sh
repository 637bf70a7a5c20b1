set -x
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/c5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c5_pytest.log
DGIBBS_VERBOSE=1 timeout 300 python bench.py --config cfg3 --nbase 2048 --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/c5_bench_cfg3.log 2>&1; echo "rc=$?" >> gpurun_out/c5_bench_cfg3.log
DGIBBS_PROFILE=1 timeout 600 python bench.py --config cfg5 --nbase 296 --nsamp 4 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c5_bench_cfg5.log 2>&1; echo "rc=$?" >> gpurun_out/c5_bench_cfg5.log
timeout 300 python bench.py --config cfg1 --nbase 1184 --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/c5_bench_cfg1.log 2>&1; echo "rc=$?" >> gpurun_out/c5_bench_cfg1.log
echo done
