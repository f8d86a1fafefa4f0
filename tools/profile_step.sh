#!/bin/bash
# ncu evidence for one benchmark step (run under gpurun, 1 GPU). Outputs land in gpurun_out/.
#   1. plain run must exit 0 first
#   2. launch list: every kernel of one timed step with its device time
#   3. ncu --set full on the top kernels (GEGLU GEMM, linear GEMM, attention, LayerNorm)
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --profile --no-library-baseline"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off \
    --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:'gemm_kernel|attn_fwd_kernel|layernorm_stream' -c 48 -o gpurun_out/prof_top $CMD > gpurun_out/ncu_full.log 2>&1
echo "full capture rc=$?"
# the report itself is too large to travel back (64 MiB limit): export the raw page here and drop it
ncu -i gpurun_out/prof_top.ncu-rep --page raw --csv > gpurun_out/prof_top_raw.csv 2> gpurun_out/ncu_export.log
echo "export rc=$?"
rm -f gpurun_out/prof_top.ncu-rep
ls -la gpurun_out/ | head -30
