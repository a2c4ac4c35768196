#!/bin/bash
# A/B of the persistent kernel's L2 eviction hints (HONK_TC2_L2HINT): conv span, step time, SM clock; then DRAM traffic by ncu.
for h in 0 1 0 1; do
HONK_TC2_L2HINT=$h timeout 300 python bench.py --steps 8 --warmup 3 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('l2hint $h: res_conv', round(d['spans_ms_per_step']['res_conv'],3), 'ms; step', round(d['ms_per_step'],3), 'ms; clocks', d['clocks']['sm_mhz'], d['clocks']['reasons'])
"
done
for h in 0 1; do
HONK_TC2_L2HINT=$h timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:tc2_res_kernel -c 1 python bench.py --steps 1 --warmup 1 2>&1 | grep -E "dram__bytes|gpu__time" | sed "s/^/l2hint $h: /"
done
