#!/bin/bash
# A/B of LBVH variant builds in one call.  usage: gpu_run_g.sh TAG variant1,variant2,... [nopytest]  (libs lib/librtb200_<variant>.so; "" = default)
T=${1:-x}; V=${2:-}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
L=$PWD/opencl-raytracing_b200/lib
for spp in 32 256; do
  timeout 200 python tools/bvh_probe.py c4,c5 $spp > $O/r02p_${T}_${spp}.log 2>&1
  for v in ${V//,/ }; do
    RTB200_LIB=$L/librtb200_$v.so timeout 200 python tools/bvh_probe.py c4,c5 $spp > $O/r02p_${T}_${v}_${spp}.log 2>&1
  done
done
[ "$3" = nopytest ] || ( time timeout 600 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_crops.py tests/test_gpu_random_scenes.py tests/test_zz_no_alloc.py -m gpu -q --timeout 300 ) > $O/r02p_${T}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02p_${T}_pytest.log
echo done
