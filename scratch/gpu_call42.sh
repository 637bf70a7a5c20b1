set -x
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29501 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/c42_bench_2gpu.log 2>&1; echo "rc=$?" >> gpurun_out/c42_bench_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > gpurun_out/c42_bench_2gpu_ref.log 2>&1; echo "rc=$?" >> gpurun_out/c42_bench_2gpu_ref.log
