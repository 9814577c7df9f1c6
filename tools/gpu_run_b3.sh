#!/bin/bash
# round-2 GPU call B3: ncu captures of the LBVH render kernel (C4, C5) and k_head, launch list of the default command
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
NCU="ncu --set full --clock-control none --import-source on"
B="--steps 1 --warmup 1 --no-cpu-baseline --no-extras"
timeout 400 $NCU -k regex:^k_trace_tail_fast_bvh\$ -s 1 -c 1 -o $O/r02b3_prof_c4 python bench.py --config c4 --spp 16 $B > $O/r02b3_ncu_c4.log 2>&1
timeout 400 $NCU -k regex:^k_trace_tail_fast_bvh\$ -s 1 -c 1 -o $O/r02b3_prof_c5 python bench.py --config c5 --spp 16 $B > $O/r02b3_ncu_c5.log 2>&1
timeout 300 $NCU -k regex:^k_head\$ -s 1 -c 1 -o $O/r02b3_prof_head python bench.py --spp 64 $B > $O/r02b3_ncu_head.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02b3_launches_c2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > $O/r02b3_launches_c2.log 2>&1
echo done
