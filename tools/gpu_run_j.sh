#!/bin/bash
# round-2 closing 1-GPU call: whole GPU suite, smoke, the default bench line (with extra_configs), the reference arm, the launch list
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
( time timeout 1200 python -m pytest tests -m gpu -q --timeout 900 -x ) > $O/r02j_pytest_all.log 2>&1; echo "pytest rc=$?" >> $O/r02j_pytest_all.log
( time timeout 120 python -c "import __graft_entry__ as g; g.smoke()" ) > $O/r02j_smoke.log 2>&1; echo "smoke rc=$?" >> $O/r02j_smoke.log
( time timeout 900 python bench.py ) > $O/r02j_bench_default.json 2> $O/r02j_bench_default.err
( time timeout 300 python bench.py --impl reference --steps 5 --warmup 3 ) > $O/r02j_bench_reference.json 2> $O/r02j_bench_reference.err
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02j_launches_c2.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras > $O/r02j_launches_c2.log 2>&1
echo done
