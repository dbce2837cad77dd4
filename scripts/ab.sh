#!/bin/bash
# A/B of environment settings on the default bench (device-timed ms/step).  Usage: bash scripts/ab.sh "ENV1=.. ENV2=.." "ENV=.." ...
# Each setting is run twice, interleaved, so that drift shows.
cd "$(dirname "$0")/.."
for rep in 1 2; do
  for e in "$@"; do
    ms=$(env $e python bench.py --steps ${STEPS:-60} --warmup 5 --no-cpu-baseline --no-roofline --no-parity ${BENCH_ARGS} 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],4))")
    echo "rep$rep [$e] $ms"
  done
done
