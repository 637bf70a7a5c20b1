set -x
timeout 900 python -m pytest tests/test_gpu_cross.py -m gpu -x -q > gpurun_out/c11_pytest_cross.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c11_pytest_cross.log
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_cross.py > gpurun_out/c11_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c11_pytest.log
python bench.py --nbase 148 --nsamp 4 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c11_bench_small.log 2>&1
echo done
