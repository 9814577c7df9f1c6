#!/bin/bash
# round-2 GPU call A: full GPU test suite, LBVH A/B variants with counted work, C2/C3 with and without the head-class sort, ncu of the C2 tail kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/r02a_smi.txt
nproc >> $O/r02a_smi.txt
( time timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -s ) > $O/r02a_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02a_pytest.log
L=$PWD/opencl-raytracing_b200/lib
python tools/bvh_probe.py c4,c5 32 > $O/r02a_probe_base.log 2>&1
for v in nostack noflag mb6; do RTB200_LIB=$L/librtb200_$v.so timeout 300 python tools/bvh_probe.py c4,c5 32 > $O/r02a_probe_$v.log 2>&1; done
for r in 4 12 16 24; do RTB200_BVH_REGEN=$r timeout 300 python tools/bvh_probe.py c4,c5 32 > $O/r02a_probe_regen$r.log 2>&1; done
timeout 600 python bench.py --steps 3 --warmup 3 > $O/r02a_bench_c2.json 2> $O/r02a_bench_c2.err
RTB200_HEAD_SORT=0 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras > $O/r02a_bench_c2_nosort.json 2>> $O/r02a_bench_c2.err
RTB200_HEAD_SORT=0x3F00 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras > $O/r02a_bench_c2_sortclass.json 2>> $O/r02a_bench_c2.err
timeout 300 python bench.py --config c3 --steps 2 --warmup 2 --no-cpu-baseline > $O/r02a_bench_c3.json 2> $O/r02a_bench_c3.err
RTB200_HEAD_SORT=0 timeout 300 python bench.py --config c3 --steps 2 --warmup 2 --no-cpu-baseline > $O/r02a_bench_c3_nosort.json 2>> $O/r02a_bench_c3.err
CMD="python bench.py --spp 256 --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_trace_tail -s 2 -c 1 -o $O/r02a_prof_c2 $CMD > $O/r02a_ncu_c2.log 2>&1
echo done
