#!/bin/bash
# One GPU session: parity tests, bench (both arms), ncu launch list of the bench command, one ncu --set full capture of a
# whole decode step (every kernel once).  Outputs land in gpurun_out/<tag>_*.
tag=${1:-r01}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem --format=csv > $out/${tag}_smi.txt 2>&1
nproc > $out/${tag}_nproc.txt; which nasm yasm >> $out/${tag}_nproc.txt 2>&1
python -m pytest tests -m gpu -x -q > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $out/${tag}_pytest_gpu.log
tail -3 $out/${tag}_pytest_gpu.log
python bench.py --steps 20 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
tail -c 600 $out/${tag}_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2> $out/${tag}_bench_ref.err; echo "ref rc=$?"
python tools/quick_bench.py 64 8 > $out/${tag}_quick.log 2>&1; echo "quick rc=$?"
grep "run 5" $out/${tag}_quick.log
python bench.py --workload encode --steps 10 --warmup 3 > $out/${tag}_bench_encode.json 2> $out/${tag}_bench_encode.err; echo "encode bench rc=$?"
tail -c 400 $out/${tag}_bench_encode.json
if [ "${SKIP_NCU:-0}" != "1" ]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --skip-cpu-baseline > $out/${tag}_ncu_list.log 2>&1; echo "ncu list rc=$?"
# one full step of the single-stream (profile) decoder after 3 warm runs: quick_bench launches 12 kernels per run
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_' --launch-skip ${NCU_SKIP:-36} --launch-count ${NCU_COUNT:-12} \
    -o $out/${tag}_full python tools/quick_bench.py 64 8 > $out/${tag}_ncu_full.log 2>&1; echo "ncu full rc=$?"
# every encoder kernel of one warm encode call on a single stream (the profile encoder of the bench: last call of the run)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'k_enc' --launch-skip ${NCU_ENC_SKIP:-390} --launch-count 10 \
    -o $out/${tag}_enc_full python bench.py --workload encode --steps 1 --warmup 3 --skip-cpu-baseline > $out/${tag}_ncu_enc_full.log 2>&1; echo "ncu enc rc=$?"
fi
ls -la $out
