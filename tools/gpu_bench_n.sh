#!/bin/bash
# gpurun --gpus N --timeout 900 -- 'bash tools/gpu_bench_n.sh N'   : both workloads on N GPUs of one box (torchrun, NCCL weight broadcast)
mkdir -p gpurun_out
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n$N.log 2>&1; echo "exit $?" >> gpurun_out/bench_n$N.log
tail -2 gpurun_out/bench_n$N.log | cut -c1-300
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --workload default_tf --batch 8192 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_tf_n$N.log 2>&1; echo "exit $?" >> gpurun_out/bench_tf_n$N.log
tail -2 gpurun_out/bench_tf_n$N.log | cut -c1-300
