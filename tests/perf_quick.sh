#!/bin/bash
# tests/perf_quick.sh [tag] -- device-resident step + per-kernel times of the four BASELINE workloads (tuning aid, run on the GPU box)
tag=${1:-q}
for wl in C3 C4 C1 C2; do
  python bench.py --no-extras --workload $wl --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['kernel_ms']
print('$wl', 'ms=%.3f'%d['ms_per_step'], d['config']['check'], ' '.join('%s=%.3f'%(a,b) for a,b in k.items()), d['config']['tiles'])
" | tee -a gpurun_out/perf_$tag.txt
done
