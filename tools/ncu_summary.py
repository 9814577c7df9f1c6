"""Summary of one ncu --set full report for profiles/: headline raw metrics + the hottest source lines.
usage: ncu_summary.py report.ncu-rep "header text" > profiles/xxx.txt"""
import csv, subprocess, sys, re
rep, header = sys.argv[1], sys.argv[2]
print(header)
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
h = rr[0]
for n in ("gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "smsp__issue_active.avg.pct_of_peak_sustained_active",
          "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
          "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
          "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
          "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
          "sm__cycles_active.avg", "sm__cycles_elapsed.avg"):
    if n in h: print("  %-70s %s %s" % (n, rr[2][h.index(n)], rr[1][h.index(n)]))
st = []
for i, n in enumerate(h):
    m = re.match(r"smsp__average_warps_issue_stalled_(.+)_per_issue_active.ratio", n)
    if m:
        try: st.append((float(rr[2][i]), m.group(1)))
        except ValueError: pass
st.sort(reverse=True)
print("  stalls per issue: " + ", ".join("%s %.2f" % (n, v) for v, n in st[:7]))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h0 = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
hh = rows[h0]
iE, iT, iS = hh.index("Instructions Executed"), hh.index("Thread Instructions Executed"), hh.index("# Samples")
per = {}; tot = 0; tot_s = 0
for r in rows[h0 + 1:]:
    if len(r) > iT and r[2] == "-" and r[0].isdigit():
        e, t, s = int(r[iE] or 0), int(r[iT] or 0), int(r[iS] or 0)
        per[int(r[0])] = (e, t, s, r[1]); tot += e; tot_s += s
print("hottest source lines of rt_kernels.cuh (share of executed warp instructions, share of stall samples, active lanes):")
for ln in sorted(per):
    e, t, s, src = per[ln]
    if 100.0 * e / tot >= 0.8:
        print("  %5d %5.2f%% smp %5.2f%% lanes %4.1f | %s" % (ln, 100.0 * e / tot, 100.0 * s / max(tot_s, 1), t / max(e, 1), src.strip()[:120]))
