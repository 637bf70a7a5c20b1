set -x
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 1 --warmup 1 --nsamp 20 --no-cpu-baseline > gpurun_out/c14_bench_2gpu.log 2>&1; echo "rc=$?" >> gpurun_out/c14_bench_2gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --impl reference --steps 1 --warmup 0 > gpurun_out/c14_bench_2gpu_ref.log 2>&1; echo "rc=$?" >> gpurun_out/c14_bench_2gpu_ref.log
echo done
