"""Executed warp instructions per code region (function) of rt_kernels.cuh, from an ncu SASS source page joined with
nvdisasm -g line info.  Regions are found from the source file itself (function definitions), math helpers are
attributed to the enclosing kernel-file line.  usage: sass_regions.py <ncu_sass.csv> <nvdisasm.sass> <kernel> [raw.csv]"""
import csv, re, sys, collections
ncu_csv, dis, kernel = sys.argv[1:4]
import os
SRC = os.environ.get("RT_SRC", "/root/repo/opencl-raytracing_b200/csrc/rt_kernels.cuh")   # the source the profiled library was built from
src = open(SRC).read().split("\n")
# region boundaries: lines starting a function definition
marks = []
for i, l in enumerate(src, 1):
    m = re.match(r"(?:__device__|inline|template).*?\b([a-z_0-9]+)\s*\(", l)
    if l.startswith("__device__") or l.startswith("inline"):
        m = re.search(r"\b([A-Za-z_0-9]+)\s*\(", l.split("__forceinline__")[-1])
        if m: marks.append((i, m.group(1)))
# finer regions inside hit_meshes_warp
for i, l in enumerate(src, 1):
    for key, name in (("if (!__any_sync(RT_FULL, need)) continue", "mesh: bound test"), ("uint32_t cnt = 0u;", "mesh: front list"),
                      ("uint32_t incl = cnt;", "mesh: scan"), ("for (uint32_t base = 0; base < total", "mesh: dispatch+MT"),
                      ("const uint32_t cand = __ballot_sync", "mesh: collect")):
        if key in l: marks.append((i, name))
marks.sort()
def region(ln):
    r = "?"
    for a, n in marks:
        if a <= ln: r = n
        else: break
    return r
rows = list(csv.reader(open(ncu_csv))); hdr = rows[1]; data = [r for r in rows[2:] if len(r) > 10]
iE, iT = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
lines = open(dis).read().split("\n")
start = [i for i, l in enumerate(lines) if l.strip().startswith(".section") and (".text." + kernel + ",") in l][0]
cur = None; last = None; k = 0; reg = collections.Counter(); regt = collections.Counter(); tot = 0
for l in lines[start + 1:]:
    if l.strip().startswith(".section"): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        if cur[0] == "rt_kernels.cuh": last = cur
        continue
    if re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l):
        if k >= len(data): break
        e, t = int(data[k][iE]), int(data[k][iT]); k += 1
        use = cur if (cur and cur[0] == "rt_kernels.cuh") else last
        r = region(use[1]) if use else "?"
        reg[r] += e; regt[r] += t; tot += e
print("kernel %s: %d SASS instructions, %.3e executed warp instructions, avg lanes %.1f" % (kernel, k, tot, sum(regt.values()) / tot))
for r, e in reg.most_common():
    print("  %-26s %5.2f%%  lanes %4.1f" % (r, 100.0 * e / tot, regt[r] / max(e, 1)))
if len(sys.argv) > 4:
    rr = list(csv.reader(open(sys.argv[4]))); h = rr[0]
    for n in ("gpu__time_duration.sum", "launch__registers_per_thread", "smsp__issue_active.avg.pct_of_peak_sustained_active",
              "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
              "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
              "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
              "dram__bytes_read.sum", "dram__bytes_write.sum"):
        if n in h: print("  %-66s %s %s" % (n, rr[2][h.index(n)], rr[1][h.index(n)]))
    stalls = []
    for i, n in enumerate(h):
        m = re.match(r"smsp__average_warps_issue_stalled_(.+)_per_issue_active.ratio", n)
        if m:
            try: stalls.append((float(rr[2][i]), m.group(1)))
            except ValueError: pass
    stalls.sort(reverse=True)
    if stalls: print("  stalls per issue: " + ", ".join("%s %.2f" % (n, v) for v, n in stalls[:6]))
    try:
        a, e = float(rr[2][h.index("sm__cycles_active.avg")]), float(rr[2][h.index("sm__cycles_elapsed.avg")])
        print("  SM active / elapsed cycles = %.3f" % (a / e))
    except (ValueError, IndexError):
        pass
