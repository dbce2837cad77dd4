#!/bin/bash
# Round-end record on one GPU: GPU tests, smoke(), full bench lines of base / big / long.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final_pytest_gpu.log; tail -4 gpurun_out/final_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; tail -2 gpurun_out/final_smoke.log
timeout 600 python bench.py --steps 100 --warmup 5 > gpurun_out/final_bench_base_n1.json 2> gpurun_out/final_bench_base_n1.err; echo "base rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference_n1.json 2>/dev/null; echo "reference rc=$?"
timeout 600 python bench.py --config big --steps 50 --warmup 5 > gpurun_out/final_bench_big_n1.json 2> gpurun_out/final_bench_big_n1.err; echo "big rc=$?"
timeout 900 python bench.py --config long --steps 10 --warmup 3 > gpurun_out/final_bench_long_n1.json 2> gpurun_out/final_bench_long_n1.err; echo "long rc=$?"
python - <<'PY'
import json
for n in ("base","big","long"):
    try:
        d=json.loads(open(f"gpurun_out/final_bench_{n}_n1.json").read().strip().splitlines()[-1])
        print(n, {k:d.get(k) for k in ("value","ms_per_step","gpu_launches","step_tflops_achieved_per_gpu","step_frac_of_gemm_roofline")}, "parity", d["parity"]["ok"], "e2e", round(d["e2e"]["value"]), "roofline", round(d["roofline"]["frac"],3), d["roofline"].get("in_step",{}).get("frac"))
    except Exception as e: print(n, "failed", e)
PY
