#!/bin/bash
# tests/sass_summary.sh -- SASS mnemonic counts and ptxas resource lines of the built library -> profiles/r02_sass_summary.txt
# (run in the build container after `make -C parallel-c-programs_b200`; needs its build.log)
cd "$(dirname "$0")/.." || exit 1
L=parallel-c-programs_b200/libb200rg.so; B=parallel-c-programs_b200/build.log
{
  echo "# SASS / ptxas evidence for libb200rg.so, build of commit $(git rev-parse --short HEAD)"
  echo "# cuobjdump -sass $L | grep -c <mnemonic>"
  cuobjdump -sass $L > /tmp/rg_sass.txt
  for m in UTMALDG UTMASTG SYNCS.ARRIVE SYNCS.PHASECHK VABSDIFF4 LDGSTS ATOMS.CAS ATOMS.OR ATOMG.E.CAS RED.E STG.E.128 'LDG.E.STRONG.SYS' 'HMMA\|IMMA\|UTCHMMA\|UTCIMMA'; do
    printf "%-28s %s\n" "$m" "$(grep -c "$m" /tmp/rg_sass.txt)"
  done
  echo
  echo "# -Xptxas -v per kernel (registers, barriers, static smem; spills)"
  python3 - "$B" <<'PY'
import re, subprocess, sys
txt = open(sys.argv[1]).read().splitlines()
out = []
for i, l in enumerate(txt):
    m = re.search(r"Compiling entry function '(\S+)' for 'sm_100a'", l)
    if not m or m.group(1).startswith("_ZN3cub"): continue
    name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
    spill = used = ""
    for k in txt[i + 1:i + 5]:
        if "bytes stack frame" in k: spill = k.strip()
        if "Used" in k: used = re.sub(r"ptxas info\s*:\s*Used ", "", k).strip(); break
    out.append("%-90s | %s | %s" % (name[:90], used, spill))
print("\n".join(out))
PY
} > profiles/r02_sass_summary.txt
wc -l profiles/r02_sass_summary.txt
