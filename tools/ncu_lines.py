"""Executed warp instructions and stall samples per SOURCE LINE of one ncu report (--import-source on).
usage: ncu_lines.py report.ncu-rep [first_line last_line] [min_pct]  — lines of rt_kernels.cuh, sorted by line number."""
import collections, csv, subprocess, sys
rep = sys.argv[1]
lo, hi = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (0, 10**9)
min_pct = float(sys.argv[4]) if len(sys.argv) > 4 else 0.3
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h0 = [i for i, r in enumerate(rows) if r and r[0] == "Line No"][0]
h = rows[h0]
iE, iT, iS = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("# Samples")
per = {}; tot = 0; tot_s = 0
for r in rows[h0 + 1:]:
    if len(r) > iT and r[2] == "-" and r[0].isdigit():
        e, t, s = int(r[iE] or 0), int(r[iT] or 0), int(r[iS] or 0)
        per[int(r[0])] = (e, t, s, r[1]); tot += e; tot_s += s
print("total warp instructions %.4e, samples %d" % (tot, tot_s))
acc = 0.0
for ln in sorted(per):
    e, t, s, src = per[ln]
    if lo <= ln <= hi: acc += 100.0 * e / tot
    if lo <= ln <= hi and 100.0 * e / tot >= min_pct:
        print("%5d %5.2f%% smp %5.2f%% lanes %4.1f | %s" % (ln, 100.0 * e / tot, 100.0 * s / max(tot_s, 1), t / max(e, 1), src.strip()[:140]))
print("lines %d..%d: %.2f%% of the instructions" % (lo, min(hi, max(per)), acc))
