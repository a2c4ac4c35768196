#!/bin/bash
# Timing experiments on the pooled models' tcgen05 pre-pass (tc2_conv0_pool_tc_kernel, csrc/res_tc2.cu): HONK_TC2_C0_MODE
# switches parts of the kernel off (bit 0 builders' loads/stores, bit 1 epilogue pooling + stores, bit 2 MMAs, bit 3 zero units
# of y0; results are garbage), HONK_TC2_CONV0_MMA=1 selects the previous mma.sync pre-pass.  Prints the res_conv0_pool span.
run() {
  timeout 200 python bench.py --model res8 --global-clips 16384 --steps 10 --warmup 3 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$1', round(d['spans_ms_per_step']['res_conv0_pool'],3), 'ms; step', round(d['ms_per_step'],3), 'ms;', round(d['value']), 'clips/s')
"
}
for m in 0 1 2 4 8 15; do HONK_TC2_C0_MODE=$m run "tcgen05 pre-pass, mode $m:"; done
HONK_TC2_CONV0_MMA=1 run "mma.sync pre-pass:"
HONK_TC2_CONV0_F32=1 run "FP32 pre-pass:"
