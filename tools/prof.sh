#!/bin/bash
# usage: tools/prof.sh <tag> [bench args]   — ncu full capture of k_trace on the GPU box + local region breakdown
TAG=$1; shift
ARGS="${@:---spp 64 --steps 2 --warmup 1 --no-cpu-baseline}"
/usr/local/graft/bin/gpurun --timeout 1500 -- "CMD=\"python bench.py $ARGS\"; \$CMD > gpurun_out/plain_$TAG.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_trace -s ${SKIP:-0} -c 1 -o gpurun_out/prof_$TAG \$CMD > gpurun_out/ncu_$TAG.log 2>&1; echo full=\$?" 2>&1 | tail -3
cd /root/repo/gpurun_out && ncu -i prof_$TAG.ncu-rep --page source --csv > /tmp/src_$TAG.csv 2>/dev/null; ncu -i prof_$TAG.ncu-rep --page raw --csv > /tmp/raw_$TAG.csv
mkdir -p /tmp/cub_$TAG && cd /tmp/cub_$TAG && rm -f *.cubin && cuobjdump -xelf all /root/repo/opencl-raytracing_b200/lib/librtb200.so > /dev/null
K=$(python -c "import csv;print(list(csv.reader(open('/tmp/src_$TAG.csv')))[0][1])")
F=rt_kernels_fast.sm_100a.cubin; [[ $K == *strict* ]] && F=rt_kernels_strict.sm_100a.cubin
nvdisasm -g $F > /tmp/cub_$TAG/k.sass
python /root/repo/tools/sass_regions.py /tmp/src_$TAG.csv /tmp/cub_$TAG/k.sass $K /tmp/raw_$TAG.csv
