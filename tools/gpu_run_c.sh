#!/bin/bash
# multi-GPU call: the real multi-process path — IPC + device barrier + pixel / spp split + shared host image.  usage: gpu_run_c.sh N
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/r02c_smi_$N.txt
if [ "$N" = "2" ]; then ( time timeout 600 python -m pytest tests/test_gpu_multiprocess.py -q --timeout 600 -s ) > $O/r02c_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02c_pytest.log; fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
B="--gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $TR 29511 bench.py $B > $O/r02c_bench_c2_${N}g_pixels.json 2> $O/r02c_${N}g_pixels.err
timeout 300 $TR 29512 bench.py $B --split spp > $O/r02c_bench_c2_${N}g_spp.json 2> $O/r02c_${N}g_spp.err

echo done
