"""Attribute executed warp instructions (ncu --page source --csv, SASS view) to CUDA source lines using
nvdisasm -g line annotations of the same cubin.  usage: sass_lines.py <ncu_sass.csv> <nvdisasm.sass> <kernel> [top]"""
import csv, re, sys, collections
ncu_csv, dis, kernel = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
rows = list(csv.reader(open(ncu_csv)))
hdr = rows[1]; data = [r for r in rows[2:] if len(r) > 10]
iS, iE, iT = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
# disassembly: find .text.<kernel> section
lines = open(dis).read().split("\n")
start = [i for i, l in enumerate(lines) if l.strip().startswith(".section") and (".text." + kernel + ",") in l][0]
cur = None; ins = []
for l in lines[start + 1:]:
    if l.strip().startswith(".section"): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        ins.append((cur, m.group(2).strip()))
print("ncu rows %d, disasm instructions %d" % (len(data), len(ins)))
n = min(len(data), len(ins))
by = collections.Counter(); byt = collections.Counter(); tot = 0
mism = 0
for k in range(n):
    e, t = int(data[k][iE]), int(data[k][iT])
    a = data[k][iS].split()[0 if not data[k][iS].startswith("@") else 1].split(".")[0]
    b = ins[k][1].split()[0 if not ins[k][1].startswith("@") else 1].split(".")[0]
    if a != b: mism += 1
    by[ins[k][0]] += e; byt[ins[k][0]] += t; tot += e
print("opcode mismatches %d" % mism)
src_cache = {}
def src(f, ln):
    if f not in src_cache:
        try: src_cache[f] = open("/root/repo/opencl-raytracing_b200/csrc/" + f).read().split("\n")
        except OSError: src_cache[f] = []
    s = src_cache[f]
    return s[ln - 1].strip()[:100] if 0 < ln <= len(s) else ""
for (key, e) in by.most_common(top):
    f, ln = key if key else ("?", 0)
    print("%5.2f%%  lanes %4.1f  %s:%d  %s" % (100.0 * e / tot, byt[key] / max(e, 1), f, ln, src(f, ln)))
