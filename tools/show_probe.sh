#!/bin/bash
# print the lines of tools/bvh_probe.py logs given as arguments
for f in "$@"; do echo "== $f"; python - "$f" <<'PY'
import json,sys
for l in open(sys.argv[1]):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    print(d["config"], d["env"], d["spp"], "Grays_eq", d["grays_equiv"], "Gq/s", d["gqueries"], "ms", d["ms"], "trace", d["ms_trace"], "nodes/q", d["nodes_per_query"], "sph/q", d["sph_per_query"])
PY
done
