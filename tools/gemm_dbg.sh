for d in 0 1 2 3 4 8 12 15; do echo "dbg=$d"; CGS_GEMM_DEBUG=$d python tools/kernel_bench.py geglu 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        r=json.loads(l); print('   ', r['case'], round(r['ms'],4), 'ms', round(r['tflops']), 'TF')
"; done
