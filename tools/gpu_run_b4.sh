#!/bin/bash
# round-2 GPU call B2: full GPU suite with the flared-ray box test, LBVH probes, the full default bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_crops.py tests/test_gpu_api_round2.py tests/test_gpu_random_scenes.py -m gpu -q --timeout 600 -s ) > $O/r02b4_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02b4_pytest.log
L=$PWD/opencl-raytracing_b200/lib
timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02b4_probe_base.log 2>&1
for v in mb6 mb8; do RTB200_LIB=$L/librtb200_$v.so timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02b4_probe_$v.log 2>&1; done
RTB200_BVH_REGEN=16 timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02b4_probe_regen16.log 2>&1
timeout 300 python bench.py --config c5 --spp 64 --steps 2 --warmup 1 --no-cpu-baseline > $O/r02b4_bench_c5.json 2> $O/r02b4_bench_c5.err
timeout 700 python bench.py --steps 3 --warmup 3 > $O/r02b4_bench_c2.json 2> $O/r02b4_bench_c2.err
echo done
