#!/bin/bash
# GEMM timing experiments (wrong results on purpose): each bit set of CGS_GEMM_EXPT is compiled into a PRIVATE copy of
# the library (never the product .so) and timed with tools/kernel_bench.py through CGS_LIB.
#   1 no GELU   2 no output stores   4 reload A instead of B   8 one MMA per k-block
# Build the variants in the container (python tools/gemm_dbg.sh build), run them on the GPU box (… run).
set -u
cd "$(dirname "$0")/.."
mkdir -p build_variants
for d in 1 2 4 8 12 15; do
  if [ "${1:-run}" = "build" ]; then
    python -c "from comfy_gen_server_b200 import build as b; b.build_variant('build_variants/libcgs_gemm_expt$d.so', ['-DCGS_GEMM_EXPT=$d'], ['geglu_gemm.cu'])"
  else
    echo "expt=$d"; CGS_LIB=build_variants/libcgs_gemm_expt$d.so python tools/kernel_bench.py geglu 2>&1 | grep '^{' | cut -c1-160
  fi
done
