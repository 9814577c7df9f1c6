#!/bin/bash
# ncu --set full captures of the LBVH render kernel on C4 and C5 (current build).  usage: gpu_run_i.sh TAG
T=${1:-x}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
NCU="ncu --set full --clock-control none --import-source on"
B="--steps 1 --warmup 1 --no-cpu-baseline --no-extras"
timeout 400 $NCU -k regex:^k_trace_tail_fast_bvh\$ -s 1 -c 1 -f -o $O/r02i_${T}_c5 python bench.py --config c5 --spp 16 $B > $O/r02i_${T}_ncu_c5.log 2>&1
timeout 400 $NCU -k regex:^k_trace_tail_fast_bvh\$ -s 1 -c 1 -f -o $O/r02i_${T}_c4 python bench.py --config c4 --spp 16 $B > $O/r02i_${T}_ncu_c4.log 2>&1
echo done
