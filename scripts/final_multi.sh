#!/bin/bash
# Round-end record on N GPUs (N = visible GPUs): multi-rank tests (2 ranks), bench lines of base (all N), big / long (8 GPUs), NVLS A/B.
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
run() {  # tag config steps [env...]
  tag=$1; cfg=$2; steps=$3; shift 3
  env "$@" timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --config $cfg --steps $steps --warmup 5 --no-roofline > gpurun_out/final_bench_${tag}_n$N.json 2> gpurun_out/final_bench_${tag}_n$N.err
  echo "== $tag N=$N rc=$?"; python - gpurun_out/final_bench_${tag}_n$N.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    a=d.get("allreduce") or {}
    print({k:d.get(k) for k in ("value","ms_per_step","n_gpus","step_tflops_achieved_per_gpu")}, "parity", d["parity"]["ok"], d["parity"].get("cross_rank_weights_identical"), "nocoll", a.get("ms_per_step_without_collectives"), "exposed_us", a.get("exposed_us"), a.get("collective","")[:20])
except Exception as e:
    print("no line:", e)
PY
}
if [ "$N" = "2" ]; then
  timeout 900 python -m pytest tests/test_data_parallel_gpu.py -m gpu -x -q > gpurun_out/final_pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final_pytest_2gpu.log; tail -3 gpurun_out/final_pytest_2gpu.log
fi
run base base 100 B200XF_X=0
if [ "$N" = "4" ]; then run base_nvls32 base 60 B200XF_NVLS=1; fi
if [ "$N" = "8" ]; then
  run base_nccl base 60 B200XF_NVLS=0
  run base_nvls12 base 60 B200XF_NVLS_CTAS=12
  run base_nvls24 base 60 B200XF_NVLS_CTAS=24
  run big big 50 B200XF_X=0
  run big_nccl big 50 B200XF_NVLS=0
  run long long 10 B200XF_X=0
fi
