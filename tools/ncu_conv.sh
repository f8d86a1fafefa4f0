# tensor-pipe utilisation and duration of the convolution kernel at the SDXL shapes, per tile-shape variant
# (CONV_VARIANT = cgs_gemm_set_variant code) and / or private library build (CGS_LIB). Run under gpurun.
for var in ${CONV_VARIANTS:-10}; do echo "VARIANT=$var LIB=${CGS_LIB:-}"; CONV_VARIANT=$var timeout 300 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:gemm_kernel --csv python tools/conv_ncu.py 2>/dev/null | grep gemm_kernel | python -c "
import sys,csv
rows=list(csv.reader(sys.stdin))
for i in range(0,len(rows),8):
    print(rows[i][4][:44], rows[i][-1], rows[i+1][-1], '| with residual', rows[i+2][-1], rows[i+3][-1])
"; done
