#!/bin/bash
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/r02i_topo_$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
B="--gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $TR 29531 bench.py $B > $O/r02i_${N}g_rank0.json 2> $O/r02i_${N}g_rank0.err
timeout 300 $TR 29532 bench.py $B --host-image shared > $O/r02i_${N}g_shared.json 2> $O/r02i_${N}g_shared.err
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras > $O/r02i_1g.json 2> $O/r02i_1g.err
echo done
