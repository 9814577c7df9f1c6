#!/bin/bash
# one ncu --set full capture of the brute-force render kernel on C2 (closing build)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
timeout 150 ncu --set full --clock-control none --import-source on -k regex:^k_trace_tail_fast\$ -s 3 -c 1 -f -o $O/r02k_c2 python bench.py --spp 512 --steps 1 --warmup 1 --no-cpu-baseline --no-extras > $O/r02k_ncu_c2.log 2>&1
echo done
