#!/bin/bash
# cumulative time of the fast labelling pass up to each RG_DEBUG_STOP point (C4 2048^3); run on the GPU box
out=gpurun_out/phase_breakdown.txt; : > $out
one() { python bench.py --no-extras --steps 5 --warmup 3 2>&1 | grep "^{" | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['roofline']['kernel_ms']['pass_a_fast'],3))" 2>/dev/null || echo fail; }
for d in 1 3 4 6 61 62 7 0; do echo "prezero stop=$d $(RG_DEBUG_STOP=$d one)" >> $out; done
for d in 1 3 7 0; do echo "noprezero stop=$d $(RG_NO_PREZERO=1 RG_DEBUG_STOP=$d one)" >> $out; done
cat $out
