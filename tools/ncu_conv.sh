for lib in "" build_variants/libcgs_gemm_expt2.so; do echo "LIB=$lib"; CGS_LIB=$lib timeout 300 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:gemm_kernel --csv python tools/conv_ncu.py 2>/dev/null | grep gemm_kernel | python -c "
import sys,csv
rows=list(csv.reader(sys.stdin))
for i in range(0,len(rows),4):
    print(rows[i][4][:40], rows[i][-1], rows[i+1][-1], '| with residual', rows[i+2][-1], rows[i+3][-1])
"; done
