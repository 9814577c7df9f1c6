#!/bin/bash
# round-2 GPU call B: full GPU suite after the non-unit-direction fix, LBVH variants, ncu captures of the MAIN launches (C2, C3, C4, C5, strict, k_head), launch list
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
( time timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -s ) > $O/r02b_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02b_pytest.log
L=$PWD/opencl-raytracing_b200/lib
python tools/bvh_probe.py c4,c5 32 > $O/r02b_probe_base.log 2>&1
for v in mb6 mb7 noflag; do RTB200_LIB=$L/librtb200_$v.so timeout 300 python tools/bvh_probe.py c4,c5 32 > $O/r02b_probe_$v.log 2>&1; done
RTB200_BVH_REGEN=16 timeout 300 python tools/bvh_probe.py c4,c5 32 > $O/r02b_probe_regen16.log 2>&1
RTB200_BVH_REGEN=16 RTB200_LIB=$L/librtb200_mb6.so timeout 300 python tools/bvh_probe.py c4,c5 32 > $O/r02b_probe_mb6_regen16.log 2>&1
timeout 900 python bench.py --steps 3 --warmup 3 > $O/r02b_bench_c2.json 2> $O/r02b_bench_c2.err
NCU="ncu --set full --clock-control none --import-source on"
B="--steps 1 --warmup 1 --no-cpu-baseline --no-extras"
timeout 600 $NCU -k regex:^k_trace_tail_fast\$ -s 3 -c 1 -o $O/r02b_prof_c2 python bench.py --spp 512 $B > $O/r02b_ncu_c2.log 2>&1
timeout 600 $NCU -k regex:^k_trace_tail_fast\$ -s 3 -c 1 -o $O/r02b_prof_c3 python bench.py --config c3 --spp 512 $B > $O/r02b_ncu_c3.log 2>&1
timeout 600 $NCU -k regex:^k_trace_tail_strict\$ -s 3 -c 1 -o $O/r02b_prof_c2_strict python bench.py --arith strict --spp 256 $B > $O/r02b_ncu_c2_strict.log 2>&1
timeout 600 $NCU -k regex:^k_trace_tail_fast_bvh\$ -s 1 -c 1 -o $O/r02b_prof_c4 python bench.py --config c4 --spp 16 $B > $O/r02b_ncu_c4.log 2>&1
timeout 600 $NCU -k regex:^k_trace_tail_fast_bvh\$ -s 1 -c 1 -o $O/r02b_prof_c5 python bench.py --config c5 --spp 16 $B > $O/r02b_ncu_c5.log 2>&1
timeout 600 $NCU -k regex:^k_head\$ -s 1 -c 1 -o $O/r02b_prof_head python bench.py --spp 64 $B > $O/r02b_ncu_head.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02b_launches_c2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > $O/r02b_launches_c2.log 2>&1
echo done
