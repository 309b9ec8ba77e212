#!/bin/bash
# One GPU session of round 2: parity tests, bench (both arms + C2), host timeline, optionally ncu.
# usage: tools/gpu_round2.sh <tag> [tests|notests] [ncu|noncu]
tag=${1:-r02a}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem --format=csv > $out/${tag}_smi.txt 2>&1
nproc > $out/${tag}_nproc.txt
if [ "${2:-tests}" = "tests" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q --durations=12 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.log
  tail -25 $out/${tag}_pytest_gpu.log
fi
timeout 600 python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
tail -c 3000 $out/${tag}_bench.json; tail -5 $out/${tag}_bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref rc=$?"
tail -c 600 $out/${tag}_bench_ref.json
timeout 600 python bench.py --workload c2 --steps 20 --warmup 3 > $out/${tag}_bench_c2.json 2> $out/${tag}_bench_c2.err; echo "c2 rc=$?"
tail -c 1500 $out/${tag}_bench_c2.json; tail -3 $out/${tag}_bench_c2.err
B200JPEG_TRACE=1 timeout 300 python tools/quick_bench.py 64 8 > $out/${tag}_quick.log 2>&1; echo "quick rc=$?"
grep -E "run 5|e2e|trace: plan|trace: wait|rounds" $out/${tag}_quick.log | tail -12
if [ "${3:-noncu}" = "ncu" ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_ncu_launches.csv \
      python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > $out/${tag}_ncu_list.log 2>&1; echo "ncu list rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_' --launch-skip ${NCU_SKIP:-36} --launch-count ${NCU_COUNT:-12} \
      -o $out/${tag}_full python tools/quick_bench.py 64 8 > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
ls -la $out | tail -20
