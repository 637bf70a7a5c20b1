set -x
timeout 900 python -m pytest tests/test_gpu_diag.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/c34_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c34_pytest.log
B="python bench.py --config cfg3 --nbase 2048 --steps 2 --warmup 2 --no-cpu-baseline"
timeout 300 $B > gpurun_out/c34_cfg3.log 2>&1; echo "rc=$?" >> gpurun_out/c34_cfg3.log
echo done
