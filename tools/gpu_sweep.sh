cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O; L=$PWD/opencl-raytracing_b200/lib
for v in mb8 mb6; do RTB200_LIB=$L/librtb200_$v.so timeout 200 python tools/bvh_probe.py c4,c5 64 > $O/r02s_$v.log 2>&1; done
timeout 200 python tools/bvh_probe.py c4,c5 64 > $O/r02s_base.log 2>&1
for r in 4 12 16; do RTB200_BVH_REGEN=$r timeout 200 python tools/bvh_probe.py c4,c5 64 > $O/r02s_regen$r.log 2>&1; done
RTB200_BVH_CHUNK_SPP=1 timeout 200 python tools/bvh_probe.py c4,c5 64 > $O/r02s_cs1.log 2>&1
