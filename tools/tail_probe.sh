#!/bin/bash
# work-list (tail) rounds: first rounds as plain full-grid launches (tail_launches) vs all rounds in the cooperative kernel
out=gpurun_out; mkdir -p $out
python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -${PYTAIL:-2}
for o in ":tail_launches=0" ":tail_launches=1" ":tail_launches=2"; do
  echo "== $o"
  B200JPEG_TRACE=1 python tools/quick_bench.py 64 8 ":fancy_upsampling=1 :profile=1 $o" 2>&1 | grep "run 5\|work-list" | tail -2 | cut -c1-330
done
bash tools/sweep.sh ":tail_launches=0" ":tail_launches=1" ":tail_launches=2" ":tail_launches=0" ":tail_launches=1" ":tail_launches=2" 2>&1 | grep -v "^ \|Traceback\|json"
