#!/bin/bash
# tests/refresh_profiles.sh -- turn the reports of tests/profile_r02.sh (gpurun_out/) into the committed profiles/r02_* files
cd "$(dirname "$0")/.." || exit 1
H=${RG_PROFILE_COMMIT:-$(git rev-parse --short HEAD)}
S="python3 tests/summarise_ncu.py"
{ echo "# Round 2 -- ncu evidence, k_pass_a_fast on C4 2048^3 (bench default)"; echo
  echo "Command (exited 0 without ncu first, same gpurun call, tests/profile_r02.sh): \`python bench.py --no-extras --steps 2 --warmup 1\`, then"
  echo "\`ncu --set full --clock-control none --import-source on -k regex:k_pass_a_fast --launch-skip 1 --launch-count 1\`. Build: commit $H."
  echo "Algorithmic bytes of this kernel per launch: 2 B x 2048^3 = 17.18 GB (reads every voxel once, clears every mask byte once)."; echo
  $S gpurun_out/r02_fast_c4.ncu-rep; } > profiles/r02_ncu_pass_a_fast_C4.md
{ echo "# Round 2 -- ncu evidence, merge / flag / pass C on C4 2048^3"; echo
  echo "Same command as r02_ncu_pass_a_fast_C4.md, \`-k regex:\"k_merge|k_pass_c|k_flag_tiles|k_rows_out\" --launch-skip 4 --launch-count 4\`. Build: commit $H."; echo
  cat profiles/notes_r02_rest_c4.txt 2>/dev/null; echo
  $S gpurun_out/r02_rest_c4.ncu-rep; } > profiles/r02_ncu_merge_flag_passc_C4.md
for wl in C3 C1 C2; do
  { echo "# Round 2 -- ncu evidence, the kernels of one grow on $wl"; echo
    echo "Command (exited 0 without ncu first): \`python bench.py --no-extras --steps 2 --warmup 1 --workload $wl\`, then \`ncu --set full --clock-control none --import-source on -k regex:\"k_pass_a_fast|k_merge|k_pass_c|k_flag_tiles|k_rows_out\" --launch-skip n --launch-count n\` (n = 5 kernels of one step; 4 for the image C2). Build: commit $H."
    cat profiles/notes_r02_$wl.txt 2>/dev/null; echo
    $S gpurun_out/r02_all_$wl.ncu-rep; } > profiles/r02_ncu_$wl.md
done
# the capture VERDICT r01 asked for by name: the kernel that labels C3's complex tiles (k_pass_a_full then, k_pass_a_fast now)
python3 - <<PY
import re
t = open("profiles/r02_ncu_C3.md").read()
sec = t[t.index("**\`void k_pass_a_fast"):t.index("**\`void k_merge")]
def val(m): return float(re.search(r"\| %s \| ([0-9.]+)" % re.escape(m), sec).group(1))
wf, bc = val("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"), val("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")
open("profiles/r02_ncu_pass_a_full_C3.md", "w").write(
    "# Round 2 -- ncu evidence for the labelling of C3's complex tiles (final build, commit $H)\n\n"
    "Round 1 labelled C3's 32768 complex tiles in k_pass_a_full (3.10 ms, after a wasted 1.00 ms attempt in the fast pass).  That kernel no longer\n"
    "runs on C3 (0.006 ms: an empty work list): the tiles are labelled inside k_pass_a_fast -- 28672 as tiles of constant rows, 4096 by the dense\n"
    "path -- so the capture asked for under this name is that kernel's.  Same report as profiles/r02_ncu_C3.md (all kernels of the step, source lines).\n"
    "Shared-memory wavefronts per conflict-free wavefront: %.0f M / (%.0f M - %.0f M) = %.2fx (the 80-byte row stride of the lean layout: two rows\n"
    "of 64 bytes at a distance of 80 share four banks).\n\n" % (wf / 1e6, wf / 1e6, bc / 1e6, wf / (wf - bc)) + sec)
PY
cp gpurun_out/r02_launches_C4_2048.csv profiles/r02_launches_C4_2048.csv
python3 - <<PY
import csv, json, subprocess, io
out = subprocess.run(["ncu", "-i", "gpurun_out/r02_fast_c4.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out))); h, u, r = rows[0], rows[1], rows[2]
def g(m):
    v = float(r[h.index(m)]); unit = u[h.index(m)]
    return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1, "ms": 1, "us": 1e-3, "s": 1e3}.get(unit, 1)
rd, wr, ms = g("dram__bytes_read.sum"), g("dram__bytes_write.sum"), g("gpu__time_duration.sum")
json.dump({"kernel": "k_pass_a_fast", "workload": "C4 2048^3, 16 seeds (bench default)", "dram_bytes_read": rd, "dram_bytes_write": wr,
           "dram_bytes_total": rd + wr, "kernel_ms": ms, "algorithmic_bytes": 2 * 2048 ** 3, "captured_at_commit": "$H",
           "source": "profiles/r02_ncu_pass_a_fast_C4.md (ncu --set full --clock-control none, one launch)"}, open("profiles/traffic_r02.json", "w"), indent=1)
print(open("profiles/traffic_r02.json").read())
PY
