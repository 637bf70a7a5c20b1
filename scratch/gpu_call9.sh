set -x
for ns in 5 20 100; do
timeout 300 python bench.py --config cfg3 --nbase 1020 --nsamp $ns --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/c9_bench_cfg3_ns$ns.log 2>&1; echo "rc=$?" >> gpurun_out/c9_bench_cfg3_ns$ns.log
done
echo done
