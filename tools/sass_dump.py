"""Dump the SASS of a kernel with executed-instruction counts and lane occupancy (ncu source page joined with nvdisasm -g).
usage: sass_dump.py <ncu_sass.csv> <nvdisasm.sass> <kernel> [min_exec_fraction]"""
import csv, re, sys
ncu_csv, dis, kernel = sys.argv[1:4]
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0
rows = list(csv.reader(open(ncu_csv))); hdr = rows[1]; data = [r for r in rows[2:] if len(r) > 10]
iS, iE, iT = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
iSt = hdr.index("Warp Stall Sampling (All Samples)")
lines = open(dis).read().split("\n")
start = [i for i, l in enumerate(lines) if l.strip().startswith(".section") and (".text." + kernel + ",") in l][0]
cur = None; k = 0
tot = sum(int(r[iE]) for r in data)
for l in lines[start + 1:]:
    if l.strip().startswith(".section"): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        if k >= len(data): break
        e, t, st = int(data[k][iE]), int(data[k][iT]), data[k][iSt]; k += 1
        if e >= thr * tot:
            print("%5d %6.3f%% l%4.1f st%5s  %-60s %s:%s" % (k, 100.0 * e / tot, t / max(e, 1), st, m.group(2)[:60], cur[0] if cur else "?", cur[1] if cur else 0))
    elif re.match(r"\s*\.L_x_\d+:", l):
        print(l.strip())
