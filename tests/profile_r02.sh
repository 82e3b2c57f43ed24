#!/bin/bash
# tests/profile_r02.sh -- the ncu captures behind profiles/r02_* (run on the GPU box, one GPU; each capture only after the
# same command has exited 0 without ncu).  Reports land in gpurun_out/; tests/summarise_ncu.py turns them into markdown.
set -x
O=gpurun_out
B="python bench.py --no-extras --steps 2 --warmup 1"
$B > $O/r02_plain_c4.json 2> $O/r02_plain_c4.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches_C4_2048.csv $B > $O/r02_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_pass_a_fast --launch-skip 1 --launch-count 1 -o $O/r02_fast_c4 $B > $O/r02_ncu_f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_merge|k_pass_c|k_flag_tiles|k_rows_out" --launch-skip 4 --launch-count 4 -o $O/r02_rest_c4 $B > $O/r02_ncu_r.log 2>&1
for wl in C3 C1 C2; do
  n=5; [ $wl = C2 ] && n=4                                       # kernels of one step that match (images have no row output kernel)
  $B --workload $wl > $O/r02_plain_$wl.json 2>> $O/r02_plain_c4.err || exit 1
  ncu --set full --clock-control none --import-source on -k regex:"k_pass_a_fast|k_merge|k_pass_c|k_flag_tiles|k_rows_out" --launch-skip $n --launch-count $n -o $O/r02_all_$wl $B --workload $wl > $O/r02_ncu_$wl.log 2>&1
done
ls -la $O/*.ncu-rep
