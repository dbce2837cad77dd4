#!/bin/bash
# All three data-parallel configurations of BASELINE.json on every visible GPU: bench lines (parity + allreduce objects) into gpurun_out/.
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
run() {  # config steps extra...
  cfg=$1; steps=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) bench.py --gpus $N --config $cfg --steps $steps --warmup 5 --no-roofline "$@" > gpurun_out/bench_${cfg}_n$N.log 2> gpurun_out/bench_${cfg}_n$N.err
  echo "== $cfg N=$N rc=$?"; python - gpurun_out/bench_${cfg}_n$N.log <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ("value","ms_per_step","n_gpus","step_tflops_achieved_per_gpu","step_frac_of_nominal_2250tf")}); print("parity", d.get("parity")); a=d.get("allreduce") or {}; print("allreduce", {k:a.get(k) for k in ("busbw_gbs_at_32MB","ms_per_step_without_collectives","exposed_us")})
except Exception as e:
    print("no line:", e)
PY
  tail -3 gpurun_out/bench_${cfg}_n$N.err
}
run base ${BASE_STEPS:-100}
run big ${BIG_STEPS:-50}
run long ${LONG_STEPS:-10}
