#!/bin/bash
# LBVH probes (counted work, flare policy variants) + the LBVH-related GPU tests
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
L=$PWD/opencl-raytracing_b200/lib
timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02f_probe_base.log 2>&1
for v in br2 br8; do RTB200_LIB=$L/librtb200_$v.so timeout 200 python tools/bvh_probe.py c5 32 > $O/r02f_probe_$v.log 2>&1; done
RTB200_BVH_CHUNK_SPP=1 timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02f_probe_cs1.log 2>&1
RTB200_BVH_CHUNK_SPP=4 timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02f_probe_cs4.log 2>&1
RTB200_BVH_CHUNK_SPP=64 timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02f_probe_cs64.log 2>&1
timeout 200 python tools/bvh_probe.py c4,c5 256 > $O/r02f_probe_256spp.log 2>&1
( time timeout 600 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_crops.py tests/test_gpu_api_round2.py tests/test_gpu_random_scenes.py tests/test_gpu_fullsize.py -m gpu -q --timeout 300 -s ) > $O/r02f_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02f_pytest.log
echo done
