#!/bin/bash
# print the bench lines of a gpu_run_h.sh call.  usage: show_h.sh TAG variant
cd /root/repo/gpurun_out
for c in c2 c3 c1 c2s; do for f in r02h_$1_$c r02h_$1_$2_$c; do python - $f.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["value"], d["unit"], "ms", d["ms_per_step"], "frac", d["roofline"]["frac"], "e2e", d["e2e"]["value"], d["clocks"]["sm_mhz"])
except Exception as e: print(sys.argv[1], "ERR", e)
PY
done; done
[ -f r02h_$1_pytest.log ] && tail -n 4 r02h_$1_pytest.log
