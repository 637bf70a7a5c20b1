set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c27_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c27_pytest.log
B="python bench.py --config cfg3 --nbase 2048 --steps 2 --warmup 2 --no-cpu-baseline"
timeout 300 $B > gpurun_out/c27_cfg3.log 2>&1; echo "rc=$?" >> gpurun_out/c27_cfg3.log
DGIBBS_NO_PROJECT_KERNEL=1 timeout 300 $B > gpurun_out/c27_cfg3_oldproj.log 2>&1
echo done
