set -x
B="python bench.py --config cfg3 --nbase 240 --steps 1 --warmup 1 --no-cpu-baseline"
timeout 300 $B > gpurun_out/c33_plain.log 2>&1; echo "rc=$?" >> gpurun_out/c33_plain.log
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"diag_resident|diag_project_fft|setup_baseline" --launch-skip 3 --launch-count 3 -o gpurun_out/r01g_diag -f $B > gpurun_out/c33_ncu.log 2>&1; echo "rc=$?" >> gpurun_out/c33_ncu.log
ls -la gpurun_out/
echo done
