#!/bin/bash
# A/B of a variant build on the brute-force configs (C2, C3, C1), then the whole GPU suite.  usage: gpu_run_h.sh TAG variant [nopytest]
T=${1:-x}; V=${2:-}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
L=$PWD/opencl-raytracing_b200/lib
B="--steps 10 --warmup 3 --no-cpu-baseline --no-extras"
for c in c2 c3 c1; do
  timeout 200 python bench.py --config $c $B > $O/r02h_${T}_$c.json 2> $O/r02h_${T}_$c.err
  [ -n "$V" ] && RTB200_LIB=$L/librtb200_$V.so timeout 200 python bench.py --config $c $B > $O/r02h_${T}_${V}_$c.json 2> $O/r02h_${T}_${V}_$c.err
done
# one rank's share of an 8-way sample split: 128-sample calls
timeout 200 python bench.py --config c2 --spp 128 $B > $O/r02h_${T}_c2s.json 2> $O/r02h_${T}_c2s.err
[ -n "$V" ] && RTB200_LIB=$L/librtb200_$V.so timeout 200 python bench.py --config c2 --spp 128 $B > $O/r02h_${T}_${V}_c2s.json 2> $O/r02h_${T}_${V}_c2s.err
[ "$3" = nopytest ] || ( time timeout 1200 python -m pytest tests -m gpu -q --timeout 900 ) > $O/r02h_${T}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02h_${T}_pytest.log
echo done
