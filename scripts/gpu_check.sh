#!/bin/bash
# scripts/gpu_check.sh — one gpurun call: GPU parity tests, bench, ncu launch list of the bench command,
# ncu full capture of the dominant kernel (level-0 blocked smoother at 512^3).
# usage: gpurun --timeout 1500 -- 'bash scripts/gpu_check.sh [tag]'
tag="${1:-r1}"
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu_$tag.log
tail -3 $out/pytest_gpu_$tag.log
python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench512 exit $?"
tail -1 $out/bench512_$tag.log | cut -c1-300
python bench.py --impl reference --steps 2 --warmup 1 > $out/bench_reference_$tag.log 2>&1; echo "bench reference exit $?"
python bench.py --steps 3 --warmup 3 --workload tgv256 --no-cpu-baseline > $out/bench256_$tag.log 2>&1; echo "bench256 exit $?"
PROF="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$PROF > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 12000 --csv --log-file $out/launches512_$tag.csv $PROF > $out/ncu_launches_$tag.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k regex:k_sweep_tbo -s 4 -c 2 -f -o $out/prof_tb512_$tag $PROF > $out/ncu_full_$tag.log 2>&1
echo "ncu full exit $?"
ncu --set full --clock-control none --import-source on -k regex:k_residual_restrict -c 1 -f -o $out/prof_rr512_$tag $PROF > $out/ncu_full_rr_$tag.log 2>&1
echo "ncu full (residual+restrict) exit $?"
ls -la $out | tail -12
