#!/bin/bash
# A/B a list of alternative builds (RG_LIB) on given workloads: bench_libs.sh "C4 C3" lib1.so lib2.so ...
wls=$1; shift
out=gpurun_out/bench_libs.txt; : > $out
for lib in "" "$@"; do for w in $wls; do
  RG_LIB=$lib python bench.py --no-extras --workload $w --steps 10 --warmup 3 2>&1 | grep "^{" | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); k=d['roofline']['kernel_ms']
print('${lib##*/}', '$w', round(d['ms_per_step'],3), {a:round(b,3) for a,b in k.items()}, d['config']['check'])" >> $out 2>&1 || echo "$lib $w fail" >> $out
done; done
cat $out
