#!/bin/bash
# work-list (tail) rounds: per-round list lengths and times (B200JPEG_TRACE), single stream and grouped
out=gpurun_out; mkdir -p $out
python -m pytest tests/test_gpu_decode.py -m gpu -x -q 2>&1 | tail -${PYTAIL:-2}
for o in ":tail_spread=1" ":tail_spread=0"; do
  echo "== $o"
  B200JPEG_TRACE=1 python tools/quick_bench.py 64 8 ":fancy_upsampling=1 :profile=1 $o" 2>&1 | grep "run 5\|work-list" | tail -3
done
echo "== grouped"
B200JPEG_TRACE=1 python tools/quick_bench.py 64 8 ":fancy_upsampling=1" 2>&1 | grep "run 5\|work-list" | tail -6
bash tools/sweep.sh ":tail_spread=1" ":tail_spread=0" 2>&1 | grep -v "^ \|Traceback\|json"
