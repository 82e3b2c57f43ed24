#!/bin/bash
# per-kernel times of the four workloads (device-resident, no extras); run on the GPU box.  $1 = output tag
out=gpurun_out/bench_all_${1:-x}.txt; : > $out
for w in C4 C3 C1 C2; do
  python bench.py --no-extras --workload $w --steps 10 --warmup 3 2>&1 | grep "^{" | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms']
print('$w', round(d['ms_per_step'],3), {a:round(b,3) for a,b in k.items()}, d['config']['check'])" >> $out 2>&1 || echo "$w fail" >> $out
done
cat $out
