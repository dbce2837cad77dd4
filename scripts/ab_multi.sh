#!/bin/bash
# A/B of environment settings on the N-GPU bench.  Usage: N=2 bash scripts/ab_multi.sh "ENV=.." "ENV=.." ...
cd "$(dirname "$0")/.."
N=${N:-2}; port=29600
for rep in $(seq 1 ${REPS:-1}); do
  for e in "$@"; do
    port=$((port + 1))
    ms=$(env $e python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py --gpus $N --config ${CONFIG:-base} --steps ${STEPS:-60} --warmup 5 --no-roofline --no-parity --no-allreduce-probe 2>gpurun_out/ab_multi.err | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],4), int(d['value']))")
    echo "rep$rep N=$N [$e] $ms"
  done
done
