set -x
python -m pytest tests -m gpu -x -q > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c2_pytest.log
DGIBBS_PROFILE=1 python bench.py --nbase 148 --nsamp 4 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c2_prof.log 2>&1; echo "rc=$?" >> gpurun_out/c2_prof.log
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c2_bench_cfg2.log 2>&1; echo "rc=$?" >> gpurun_out/c2_bench_cfg2.log
python bench.py --config cfg3 --nbase 2048 --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/c2_bench_cfg3.log 2>&1; echo "rc=$?" >> gpurun_out/c2_bench_cfg3.log
echo done
