#!/bin/bash
# One GPU session of round 2: parity tests, bench (both arms, C2, encode), host timeline, ncu launch list + full capture,
# compute-sanitizer logs.   usage: tools/gpu_round2.sh <tag> [tests|notests] [ncu|noncu] [san|nosan]
tag=${1:-r02a}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem --format=csv > $out/${tag}_smi.txt 2>&1
nproc > $out/${tag}_nproc.txt
if [ "${2:-tests}" = "tests" ]; then
  rm -f $out/configs_bench.jsonl
  timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.log
  tail -14 $out/${tag}_pytest_gpu.log
  cp $out/configs_bench.jsonl $out/${tag}_configs_bench.jsonl 2>/dev/null
fi
timeout 600 python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
tail -c 600 $out/${tag}_bench.json; tail -3 $out/${tag}_bench.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --workload c2 --steps 20 --warmup 3 > $out/${tag}_bench_c2.json 2> $out/${tag}_bench_c2.err; echo "c2 rc=$?"
timeout 600 python bench.py --workload encode --steps 10 --warmup 3 > $out/${tag}_bench_encode.json 2> $out/${tag}_bench_encode.err; echo "encode rc=$?"
B200JPEG_TRACE=1 timeout 300 python tools/quick_bench.py 64 8 > $out/${tag}_stage_times.log 2>&1; echo "quick rc=$?"
grep -E "run 5|trace: plan" $out/${tag}_stage_times.log | tail -3 | cut -c1-250
timeout 600 python tools/prog_probe.py > $out/${tag}_progressive_probe.log 2>&1; cat $out/${tag}_progressive_probe.log
if [ "${3:-noncu}" = "ncu" ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_ncu_launches.csv \
      python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > $out/${tag}_ncu_list.log 2>&1; echo "ncu list rc=$?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_' --launch-skip ${NCU_SKIP:-36} --launch-count ${NCU_COUNT:-12} \
      -o $out/${tag}_full python tools/quick_bench.py 64 8 > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
if [ "${4:-nosan}" = "san" ]; then # (compute-sanitizer is closed on this pool: profiles/r02t_sanitizer_unavailable.txt)
  timeout 900 compute-sanitizer --tool memcheck --log-file $out/${tag}_sanitizer_memcheck.log python -m pytest tests/test_gpu_decode.py -x -q -k "mixed_batch or bad_and or damaged" > $out/${tag}_sanitizer_memcheck.out 2>&1; echo "memcheck rc=$?"
  tail -3 $out/${tag}_sanitizer_memcheck.log
  timeout 900 compute-sanitizer --tool racecheck --log-file $out/${tag}_sanitizer_racecheck.log python -m pytest tests/test_gpu_decode.py -x -q -k "mixed_batch" > $out/${tag}_sanitizer_racecheck.out 2>&1; echo "racecheck rc=$?"
  tail -3 $out/${tag}_sanitizer_racecheck.log
fi
ls -la $out | tail -30
