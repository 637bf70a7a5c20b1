set -x
DGIBBS_VERBOSE=1 timeout 900 python -m pytest tests/test_gpu_diag.py -m gpu -x -q -k variants -s > gpurun_out/c40_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c40_pytest.log
