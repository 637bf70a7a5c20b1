set -x
A3="--config cfg3 --nbase 1020 --steps 1 --warmup 1 --no-cpu-baseline"
python bench.py $A3 > gpurun_out/c10_plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:diag_resident -s 1 -c 1 -o gpurun_out/c10_diag_full python bench.py $A3 > gpurun_out/c10_ncu3.log 2>&1
echo done
