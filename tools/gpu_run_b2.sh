#!/bin/bash
# round-2 GPU call B2: full GPU suite with the flared-ray box test, LBVH probes, the full default bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
( time timeout 900 python -m pytest tests -m gpu -q --timeout 600 -s ) > $O/r02b2_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02b2_pytest.log
L=$PWD/opencl-raytracing_b200/lib
timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02b2_probe_base.log 2>&1
for v in mb6 mb8; do RTB200_LIB=$L/librtb200_$v.so timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02b2_probe_$v.log 2>&1; done
RTB200_BVH_REGEN=16 timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02b2_probe_regen16.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 > $O/r02b2_bench_c2.json 2> $O/r02b2_bench_c2.err
echo done
