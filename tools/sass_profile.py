"""Summarise an `ncu --page source --csv` dump: executed warp instructions per opcode and lane utilisation."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; data = rows[2:]
iS, iE, iT = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
iSamp = hdr.index("# Samples")
by_op = collections.Counter(); thr_by_op = collections.Counter(); tot = 0; tott = 0
for r in data:
    try: e = int(r[iE]); t = int(r[iT])
    except (ValueError, IndexError): continue
    op = r[iS].split()
    if not op: continue
    o = op[1] if op[0].startswith("@") else op[0]
    o = o.split(".")[0]
    by_op[o] += e; thr_by_op[o] += t; tot += e; tott += t
print("total warp inst %d, thread inst %d, avg lanes %.2f" % (tot, tott, tott / tot))
for o, e in by_op.most_common(28):
    print("%-10s %6.2f%%  lanes %.1f" % (o, 100.0 * e / tot, thr_by_op[o] / max(e, 1)))
if len(sys.argv) > 2:
    n = int(sys.argv[2])
    # hottest contiguous windows of 40 instructions
    ex = []
    for r in data:
        try: ex.append(int(r[iE]))
        except (ValueError, IndexError): ex.append(0)
    import itertools
    win = 40
    sums = [sum(ex[i:i+win]) for i in range(0, len(ex), win)]
    for i in sorted(range(len(sums)), key=lambda i: -sums[i])[:n]:
        print("window @%d: %.2f%% of executed" % (i * win, 100.0 * sums[i] / tot))
