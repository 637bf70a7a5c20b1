set -x
B="python bench.py --config cfg3 --nbase 2048 --steps 2 --warmup 2 --no-cpu-baseline"
timeout 300 $B > gpurun_out/c23_plain.log 2>&1; echo "rc=$?" >> gpurun_out/c23_plain.log
DGIBBS_PROFILE=1 timeout 300 $B > gpurun_out/c23_prof.log 2>&1; echo "rc=$?" >> gpurun_out/c23_prof.log
DGIBBS_DIAG_HELPERS=1 timeout 300 $B > gpurun_out/c23_helpers.log 2>&1; echo "rc=$?" >> gpurun_out/c23_helpers.log
DGIBBS_DIAG_HELPERS=1 DGIBBS_PROFILE=1 timeout 300 $B > gpurun_out/c23_helpers_prof.log 2>&1; echo "rc=$?" >> gpurun_out/c23_helpers_prof.log
echo done
