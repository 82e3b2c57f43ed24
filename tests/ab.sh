#!/bin/bash
# tests/ab.sh <lib_a.so> <lib_b.so> [workloads...] -- the same workloads with two builds of the library on ONE box, alternating (tuning aid)
A=$1; B=$2; shift 2
W=${@:-C4 C3}
for rep in 1 2; do for wl in $W; do for L in $A $B; do
  RG_LIB=$PWD/$L python bench.py --no-extras --workload $wl --steps 8 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms']
print('$L', '$wl', 'ms=%.3f'%d['ms_per_step'], ' '.join('%s=%.3f'%(a,b) for a,b in k.items()))"
done; done; done
