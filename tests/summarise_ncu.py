#!/usr/bin/env python3
"""tests/summarise_ncu.py <report.ncu-rep> [ntiles] -- markdown summary of an ncu report (raw metrics per kernel launch + the
source lines that carry the most stall samples / instructions).  Run where `ncu` is installed (the build container)."""
import csv, io, subprocess, sys

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
           "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers", "launch__grid_size", "launch__block_size",
           "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
           "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
           "smsp__inst_executed_op_shared_atom.sum", "lts__t_sector_hit_rate.pct", "dram__throughput.avg.pct_of_peak_sustained_elapsed"]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, u = rows[0], rows[1]
    res = []
    for r in rows[2:]:
        d = {"kernel": r[h.index("Kernel Name")]}
        for m in METRICS:
            if m in h:
                d[m] = (r[h.index(m)], u[h.index(m)])
        res.append(d)
    return res


def source(rep, top=12):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    kernels, cur_k, cur_f = {}, None, None
    for r in csv.reader(io.StringIO(out)):
        if not r:
            continue
        if r[0] == "File Path":
            cur_f = r[1].split("/")[-1]; continue
        if r[0] == "Function Name":
            cur_k = r[1]; kernels.setdefault(cur_k, {}); continue
        if r[0] in ("Line No",) or r[0] == "":
            continue
        try:
            key = (cur_f, int(r[0]), r[1].strip()[:110])
            a = kernels[cur_k].setdefault(key, [0, 0])
            a[0] += int(r[4] or 0); a[1] += int(r[7] or 0)
        except Exception:
            pass
    res = {}
    for k, lines in kernels.items():
        ts = sum(v[0] for v in lines.values()) or 1; ti = sum(v[1] for v in lines.values()) or 1
        best = sorted(lines.items(), key=lambda kv: -(kv[1][0] / ts + kv[1][1] / ti))[:top]
        res[k] = (ts, ti, [(f, ln, src, 100 * v[0] / ts, 100 * v[1] / ti) for (f, ln, src), v in best])
    return res


def main():
    rep = sys.argv[1]
    print(f"### `{rep.split('/')[-1]}`\n")
    for d in raw(rep):
        print(f"**`{d['kernel'][:100]}`**\n\n| metric | value | unit |\n|---|---|---|")
        for m in METRICS:
            if m in d:
                print(f"| {m} | {d[m][0]} | {d[m][1]} |")
        print()
    for k, (ts, ti, lines) in source(rep).items():
        print(f"Hottest source lines of `{k[:80]}` (share of warp-stall samples / of executed warp instructions; {ti} warp instructions in all):\n")
        print("| file:line | samples % | inst % | source |\n|---|---|---|---|")
        for f, ln, src, ps, pi in lines:
            print(f"| {f}:{ln} | {ps:.1f} | {pi:.1f} | `{src.replace('|', '/')}` |")
        print()


if __name__ == "__main__":
    main()
