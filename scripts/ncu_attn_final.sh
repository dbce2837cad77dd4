#!/bin/bash
# ncu of the attention kernels in their final round-2 form (1 GPU).  Each profiled command first runs WITHOUT ncu and must exit 0.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_tensor.sum,smsp__inst_executed_pipe_xu.sum
python scripts/kernel_bench.py long --once > gpurun_out/kb_long_plain.log 2>&1 || { echo "kernel_bench long --once failed"; tail -5 gpurun_out/kb_long_plain.log; exit 1; }
ncu --metrics $M --clock-control none --csv -k regex:attn --log-file gpurun_out/ncu_attn_long_final.csv python scripts/kernel_bench.py long --once > gpurun_out/ncu_attn_long.log 2>&1
echo "ncu attn long rc=$?"
ncu --set full --clock-control none --import-source on -k regex:attn_fwd_tc_pp -s 1 -c 1 -o gpurun_out/attn_fwd_pp_final -f python scripts/kernel_bench.py long --once > gpurun_out/ncu_full_attn.log 2>&1
echo "ncu full pp rc=$?"
ncu --set full --clock-control none --import-source on -k regex:attn_bwd_tc_multi -s 1 -c 1 -o gpurun_out/attn_bwd_multi_final -f python scripts/kernel_bench.py long --once > gpurun_out/ncu_full_attn_bwd.log 2>&1
echo "ncu full bwd rc=$?"
python scripts/kernel_bench.py --once > gpurun_out/kb_once_plain.log 2>&1 && \
ncu --metrics $M --clock-control none --csv -k regex:attn --log-file gpurun_out/ncu_attn_base_final.csv python scripts/kernel_bench.py --once > gpurun_out/ncu_attn_base.log 2>&1
echo "ncu attn base rc=$?"
ls -la gpurun_out/*final*
