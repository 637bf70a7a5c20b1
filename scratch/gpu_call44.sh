set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c44_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c44_pytest.log
timeout 1500 python bench.py > gpurun_out/c44_bench_default.log 2>&1; echo "rc=$?" >> gpurun_out/c44_bench_default.log
