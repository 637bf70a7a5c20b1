set -x
timeout 900 python scratch/perf_extra.py > gpurun_out/c18_perf_extra.log 2>&1; echo "rc=$?" >> gpurun_out/c18_perf_extra.log
echo done
