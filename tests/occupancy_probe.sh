#!/bin/bash
# how the fast labelling pass reacts to fewer CTAs per SM (extra dynamic shared memory): 3 (default), 2, 1
out=gpurun_out/occupancy_probe.txt; : > $out
for pad in ${PADS:-0 40000 120000}; do
  RG_FAST_PAD_SMEM=$pad python bench.py --no-extras --steps 5 --warmup 3 2>&1 | grep "^{" | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('pad $pad', round(d['roofline']['kernel_ms']['pass_a_fast'],3), d['config']['check'])" >> $out 2>&1 || echo "pad $pad fail" >> $out
done
cat $out
