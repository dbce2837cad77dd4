#!/bin/bash
# GPU-box check of the whole tree: GPU tests, smoke(), one bench line.  Logs land in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
timeout 1300 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps ${BENCH_STEPS:-100} --warmup 5 > gpurun_out/bench.log 2> gpurun_out/bench.err; tail -c 6000 gpurun_out/bench.log; tail -5 gpurun_out/bench.err
