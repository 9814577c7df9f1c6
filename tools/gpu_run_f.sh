#!/bin/bash
# the driver's own command at N GPUs (default arguments, extra_configs included) + the reference arm
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
( time timeout 900 $TR 29541 bench.py --gpus $N ) > $O/r02f_default_${N}g.json 2> $O/r02f_default_${N}g.err
echo done
