set -x
timeout 900 python -m pytest tests/test_gpu_nrml.py -m gpu -x -q > gpurun_out/c38_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c38_pytest.log
timeout 1500 python scratch/perf_nrml.py > gpurun_out/c38_perf_nrml.log 2>&1; echo rc=$? >> gpurun_out/c38_perf_nrml.log
