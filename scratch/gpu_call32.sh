set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/c32_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c32_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/c32_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/c32_smoke.log
timeout 1500 python bench.py > gpurun_out/c32_bench_default.log 2>&1; echo "rc=$?" >> gpurun_out/c32_bench_default.log
timeout 900 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/c32_bench_reference.log 2>&1; echo "rc=$?" >> gpurun_out/c32_bench_reference.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c32_launches.csv python bench.py --nbase 296 --nsamp 10 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/c32_ncu_bench.log 2>&1; echo "rc=$?" >> gpurun_out/c32_ncu_bench.log
echo done
