set -x
B="python bench.py --config cfg3 --nbase 2048 --steps 2 --warmup 2 --no-cpu-baseline"
timeout 300 $B > gpurun_out/c24_plain.log 2>&1; echo "rc=$?" >> gpurun_out/c24_plain.log
DGIBBS_PROFILE=1 timeout 300 $B > gpurun_out/c24_prof.log 2>&1; echo "rc=$?" >> gpurun_out/c24_prof.log
echo done
