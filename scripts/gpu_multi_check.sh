#!/bin/bash
# N-GPU check (N = visible GPUs): the multi-rank parity tests, then the N-GPU bench line (parity + allreduce objects).
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_data_parallel_gpu.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi.log
tail -8 gpurun_out/pytest_multi.log
for n in ${NS:-$N}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --steps ${BENCH_STEPS:-50} --warmup 5 ${BENCH_ARGS} > gpurun_out/bench_n$n.log 2> gpurun_out/bench_n$n.err
  echo "bench N=$n rc=$?"; tail -c 3500 gpurun_out/bench_n$n.log; tail -3 gpurun_out/bench_n$n.err
done
