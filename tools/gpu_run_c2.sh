#!/bin/bash
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
B="--gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 $TR 29521 bench.py $B > $O/r02h_${N}g_rank0.json 2> $O/r02h_${N}g_rank0.err
timeout 300 $TR 29522 bench.py $B --host-image shared > $O/r02h_${N}g_shared.json 2> $O/r02h_${N}g_shared.err
timeout 300 $TR 29523 bench.py $B --reduce fused_nccl_barriers > $O/r02h_${N}g_ncclbar.json 2> $O/r02h_${N}g_ncclbar.err
timeout 300 $TR 29524 bench.py $B --reduce nccl > $O/r02h_${N}g_nccl.json 2> $O/r02h_${N}g_nccl.err
echo done
