set -x
timeout 900 python -m pytest tests/test_gpu_task.py -m gpu -x -q > gpurun_out/c36_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c36_pytest.log
echo done
