#!/bin/bash
# Runs the GPU test groups in separate processes (a sticky CUDA fault in one group must not hide the others).
# Usage (on the GPU box): bash scripts/gpu_check.sh [group ...]
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
groups=("$@")
[ ${#groups[@]} -eq 0 ] && groups=(gemm attention layernorm embedding positional vocab_ce adam train_step)
for g in "${groups[@]}"; do
  case $g in
    train_step) sel="tests/test_train_step_gpu.py" ;;
    *) sel="tests/test_kernels_gpu.py -k $g" ;;
  esac
  echo "=== $g"
  timeout 600 python -m pytest $sel -q -m gpu --timeout 120 --timeout-method=thread $XFLAG -q 2>&1 | grep -v "Warning\|warnings.html\|^$\|Triggered\|Consider" | tail -${TAILN:-40} | tee gpurun_out/test_$g.log
done
