set -x
A3="--config cfg3 --nbase 1024 --steps 1 --warmup 1 --no-cpu-baseline"
A2="--nbase 148 --nsamp 2 --steps 1 --warmup 1 --no-cpu-baseline"
A2L="--nbase 296 --nsamp 10 --steps 2 --warmup 1 --no-cpu-baseline"
python bench.py $A3 > gpurun_out/c6_plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:diag_resident -s 1 -c 1 -o gpurun_out/c6_diag_full python bench.py $A3 > gpurun_out/c6_ncu3.log 2>&1
python bench.py $A2 > gpurun_out/c6_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:dense_sampler -s 2 -c 1 -o gpurun_out/c6_dense_full python bench.py $A2 > gpurun_out/c6_ncu2.log 2>&1
python bench.py $A2L > gpurun_out/c6_plain2l.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c6_launches.csv python bench.py $A2L > gpurun_out/c6_ncu2l.log 2>&1
echo done
