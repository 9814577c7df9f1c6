#!/bin/bash
# round-2 GPU call C (2 GPUs): the real multi-process path — IPC + device barrier + pixel split + shared host image
cd "$GRAFT_REPO_ROOT" 2>/dev/null || cd /root/repo
O=gpurun_out
nvidia-smi -L > $O/r02c_smi.txt
( time timeout 900 python -m pytest tests/test_gpu_multiprocess.py tests/test_gpu_api_round2.py -q --timeout 600 -s ) > $O/r02c_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02c_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
B="--gpus 2 --steps 5 --warmup 3 --no-cpu-baseline --no-extras"
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras > $O/r02c_bench_c2_1g.json 2> $O/r02c_bench_1g.err
timeout 300 $TR 29511 bench.py $B > $O/r02c_bench_c2_2g_pixels.json 2> $O/r02c_2g_pixels.err
timeout 300 $TR 29512 bench.py $B --split spp > $O/r02c_bench_c2_2g_spp.json 2> $O/r02c_2g_spp.err
timeout 300 $TR 29513 bench.py $B --split spp --reduce fused_nccl_barriers > $O/r02c_bench_c2_2g_spp_ncclbar.json 2> $O/r02c_2g_spp_ncclbar.err
timeout 300 $TR 29514 bench.py $B --reduce nccl > $O/r02c_bench_c2_2g_pixels_nccl.json 2> $O/r02c_2g_pixels_nccl.err
timeout 600 $TR 29515 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > $O/r02c_bench_c2_2g_full.json 2> $O/r02c_2g_full.err
echo done
