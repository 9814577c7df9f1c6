#!/bin/bash
# round-2 final 1-GPU call: whole GPU suite, the default bench line (with extra_configs), ncu captures of the LBVH kernel + k_head, launch list
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
( time timeout 1200 python -m pytest tests -m gpu -q --timeout 900 -x ) > $O/r02e_pytest_all.log 2>&1; echo "pytest rc=$?" >> $O/r02e_pytest_all.log
( time timeout 900 python bench.py ) > $O/r02e_bench_default.json 2> $O/r02e_bench_default.err
( time timeout 300 python bench.py --impl reference --steps 5 --warmup 3 ) > $O/r02e_bench_reference.json 2> $O/r02e_bench_reference.err
( time timeout 120 python -c "import __graft_entry__ as g; g.smoke()" ) > $O/r02e_smoke.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
B="--steps 1 --warmup 1 --no-cpu-baseline --no-extras"
timeout 400 $NCU -k regex:^k_trace_tail_fast_bvh\$ -s 1 -c 1 -o $O/r02e_prof_c4 python bench.py --config c4 --spp 16 $B > $O/r02e_ncu_c4.log 2>&1
timeout 400 $NCU -k regex:^k_trace_tail_fast_bvh\$ -s 1 -c 1 -o $O/r02e_prof_c5 python bench.py --config c5 --spp 16 $B > $O/r02e_ncu_c5.log 2>&1
timeout 300 $NCU -k regex:^k_head\$ -s 1 -c 1 -o $O/r02e_prof_head python bench.py --spp 64 $B > $O/r02e_ncu_head.log 2>&1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02e_launches_c2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > $O/r02e_launches_c2.log 2>&1
echo done
