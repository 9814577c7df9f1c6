#!/bin/bash
# LBVH probes (counted work) + the LBVH-related GPU tests.  usage: gpu_run_d.sh TAG
T=${1:-x}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
timeout 200 python tools/bvh_probe.py c4,c5 32 > $O/r02p_${T}_32.log 2>&1
timeout 200 python tools/bvh_probe.py c4,c5 256 > $O/r02p_${T}_256.log 2>&1
( time timeout 600 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_crops.py tests/test_gpu_api_round2.py tests/test_gpu_random_scenes.py -m gpu -q --timeout 300 ) > $O/r02p_${T}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02p_${T}_pytest.log
echo done
