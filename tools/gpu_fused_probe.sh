#!/bin/bash
# A/B of the fused entropy-decode kernel against the separate kernels: stage times of 64 1080p images, then the GPU tests.
tag=${1:-r02r}
out=gpurun_out
mkdir -p $out
timeout 300 python tools/quick_bench.py 64 8 ":fancy_upsampling=1 :profile=1" > $out/${tag}_stage_times_fused.log 2>&1; echo "fused rc=$?"
grep -E "run [45]" $out/${tag}_stage_times_fused.log | cut -c1-420
timeout 300 python tools/quick_bench.py 64 8 ":fancy_upsampling=1 :profile=1 :fused=0" > $out/${tag}_stage_times_separate.log 2>&1; echo "separate rc=$?"
grep -E "run [45]" $out/${tag}_stage_times_separate.log | cut -c1-420
timeout 300 python tools/quick_bench.py 64 8 ":fancy_upsampling=1" > $out/${tag}_groups_fused.log 2>&1
grep -E "run [45]" $out/${tag}_groups_fused.log | cut -c1-200
timeout 300 python tools/quick_bench.py 64 8 ":fancy_upsampling=1 :fused=0" > $out/${tag}_groups_separate.log 2>&1
grep -E "run [45]" $out/${tag}_groups_separate.log | cut -c1-200
if [ "${2:-tests}" = "tests" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.log
  tail -12 $out/${tag}_pytest_gpu.log
fi
