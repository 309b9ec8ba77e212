#!/bin/bash
# ncu --set full of one kernel (regex $2) of the 64 x 1080p probe; tag $1; options $3
tag=${1:-r02r}; k=${2:-k_huff_fused}; opts=${3:-":fancy_upsampling=1 :profile=1"}
out=gpurun_out; mkdir -p $out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$k" --launch-skip ${NCU_SKIP:-3} --launch-count ${NCU_COUNT:-1} \
   -o $out/${tag}_${k}_full python tools/quick_bench.py 64 8 "$opts" > $out/${tag}_${k}_ncu.log 2>&1; echo "ncu rc=$?"
tail -3 $out/${tag}_${k}_ncu.log
