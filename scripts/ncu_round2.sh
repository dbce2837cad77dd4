#!/bin/bash
# ncu evidence of round 2 (1 GPU).  Each profiled command first runs WITHOUT ncu and must exit 0.  Outputs under gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
# 1. launch list of one Transformer-base train step (per-kernel gpu__time_duration)
bash scripts/ncu_launch_list.sh r02 2>&1 | tail -3
# 2. DRAM traffic + duration of the HBM-bound kernels and the attention kernels (base shapes), each launched twice (cold, warm)
python scripts/kernel_bench.py --once > gpurun_out/kb_once_plain.log 2>&1 || { echo "kernel_bench --once failed"; exit 1; }
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_tensor.sum \
    --clock-control none --csv --log-file gpurun_out/ncu_small_kernels_r02.csv python scripts/kernel_bench.py --once > gpurun_out/ncu_small.log 2>&1
echo "ncu small rc=$?"
# 3. the long-sequence attention kernels (B8 H8 S1024)
python scripts/kernel_bench.py long --once > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_tensor.sum,smsp__inst_executed_pipe_xu.sum \
    --clock-control none --csv -k regex:attn --log-file gpurun_out/ncu_attn_long_r02.csv python scripts/kernel_bench.py long --once > gpurun_out/ncu_attn.log 2>&1
echo "ncu attn rc=$?"
# 4. ncu --set full of the step's dominant GEMM launch (vocabulary projection, CTA pairs) and of the long forward attention kernel
python scripts/gemm_one.py > gpurun_out/gemm_one_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:gemm_kernel -s 6 -c 1 -o gpurun_out/gemm_vocab_r02 -f python scripts/gemm_one.py > gpurun_out/ncu_full_gemm.log 2>&1
echo "ncu full gemm rc=$?"
ncu --set full --clock-control none --import-source on -k regex:attn_fwd_tc_pp -s 1 -c 1 -o gpurun_out/attn_fwd_pp_r02 -f python scripts/kernel_bench.py long --once > gpurun_out/ncu_full_attn.log 2>&1
echo "ncu full attn rc=$?"
ls -la gpurun_out/*.ncu-rep
