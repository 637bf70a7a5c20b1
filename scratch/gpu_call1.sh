set -x
python -m pytest tests -m gpu -x -q > gpurun_out/c1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
python bench.py --steps 2 --warmup 3 > gpurun_out/c1_bench_cfg2.log 2>&1; echo "rc=$?" >> gpurun_out/c1_bench_cfg2.log
python bench.py --config cfg3 --nbase 2048 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c1_bench_cfg3.log 2>&1; echo "rc=$?" >> gpurun_out/c1_bench_cfg3.log
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/c1_bench_ref.log 2>&1; echo "rc=$?" >> gpurun_out/c1_bench_ref.log
python bench.py --nbase 148 --nsamp 2 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c1_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:dense_sampler -s 2 -c 1 -o gpurun_out/c1_dense_full python bench.py --nbase 148 --nsamp 2 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c1_ncu.log 2>&1
echo done
