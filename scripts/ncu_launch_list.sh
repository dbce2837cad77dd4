#!/bin/bash
# ncu launch list of one train step (per-kernel gpu__time_duration), written to gpurun_out/launches_<tag>.csv.
# The same command runs first WITHOUT ncu and must exit 0.  B200XF_GRAPH=0 keeps the step on stream launches so that
# ncu sees the individual kernels (the product path replays the same kernels from a CUDA graph).
cd "$(dirname "$0")/.."
tag=${1:-r01}
export B200XF_GRAPH=0
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-roofline --no-parity > gpurun_out/plain_$tag.json 2> gpurun_out/plain_$tag.err || { echo "plain run failed"; exit 1; }
L=$(python -c "import json; d=json.loads(open('gpurun_out/plain_$tag.json').read().strip().splitlines()[-1]); print(d['gpu_launches'])")
echo "launches per step: $L"
# bench.py: 3 warm-up steps, then the timed step; the e2e leg follows.  Skip the warm-up, keep one step (+ a few).
ncu --metrics gpu__time_duration.sum --clock-control none -s $((3 * L)) -c $((L + 4)) --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-roofline --no-parity > gpurun_out/ncu_$tag.log 2>&1
echo "ncu rc=$?"; wc -l gpurun_out/launches_$tag.csv
