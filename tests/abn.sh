#!/bin/bash
# tests/abn.sh <workload> <lib.so>... -- one workload with several builds of the library on ONE box, round-robin twice (tuning aid)
W=$1; shift
for rep in 1 2; do for L in "$@"; do
  RG_LIB=$PWD/$L python bench.py --no-extras --workload $W --steps 8 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms']
print('$L', '$W', 'ms=%.3f'%d['ms_per_step'], ' '.join('%s=%.3f'%(a,b) for a,b in k.items()))"
done; done
