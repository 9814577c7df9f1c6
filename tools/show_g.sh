#!/bin/bash
# print the probe lines of a gpu_run_g.sh call.  usage: show_g.sh TAG variant1,variant2
cd /root/repo/gpurun_out
for spp in 32 256; do for f in r02p_$1_$spp.log $(for v in ${2//,/ }; do echo r02p_$1_${v}_$spp.log; done); do python - $f <<'PY'
import json,sys
for l in open(sys.argv[1]):
    try: d=json.loads(l)
    except Exception: print(sys.argv[1], l.strip()[:200]); continue
    print("%-28s %s spp %3d Grays %.3f Gq %.3f ms %8.2f nodes/q %.2f sph/q %.2f tri/q %.2f" % (d["lib"], d["config"], d["spp"], d["grays_equiv"], d["gqueries"], d["ms"], d["nodes_per_query"], d["sph_per_query"], d["tri_per_query"]))
PY
done; done
[ -f r02p_$1_pytest.log ] && tail -n 5 r02p_$1_pytest.log
