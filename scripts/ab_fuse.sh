#!/bin/bash
# A/B of the fused-epilogue switches on one configuration: bash scripts/ab_fuse.sh <config> <steps> [extra bench args]
cfg=${1:-base}; steps=${2:-100}; shift 2
for v in "B200XF_FUSE_LN=0 B200XF_FUSE_COLSUM=0" "B200XF_FUSE_LN=1 B200XF_FUSE_COLSUM=0" "B200XF_FUSE_LN=0 B200XF_FUSE_COLSUM=1" "B200XF_FUSE_LN=1 B200XF_FUSE_COLSUM=1"; do
  echo -n "$cfg [$v] "
  env $v python bench.py --config $cfg --steps $steps --warmup 5 --no-cpu-baseline --no-roofline --no-parity "$@" 2>&1 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms/step', round(d['ms_per_step'],4), 'launches/step', d['gpu_launches']//d['steps'], 'loss', d['loss'])"
done
