#!/bin/bash
# scripts/gpu_check.sh — one gpurun call: GPU parity tests, bench, ncu launch list, ncu full capture.
# usage: gpurun --timeout 1500 -- 'bash scripts/gpu_check.sh [tag] [kernel-regex]'
tag="${1:-r1}"
kre="${2:-k_sweep}"
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv -lms 500 > $out/clocks_$tag.csv &
smi=$!
python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu_$tag.log
tail -3 $out/pytest_gpu_$tag.log
python bench.py --steps 3 --warmup 3 --workload tgv512 > $out/bench512_$tag.log 2>&1; echo "bench512 exit $?"
tail -1 $out/bench512_$tag.log
python bench.py --steps 3 --warmup 3 --workload tgv256 --no-cpu-baseline > $out/bench256_$tag.log 2>&1; echo "bench256 exit $?"
kill $smi
PROF="python bench.py --steps 1 --warmup 1 --workload tgv256 --no-e2e --no-cpu-baseline"
$PROF > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $out/launches_$tag.csv $PROF > $out/ncu_launches_$tag.log 2>&1
echo "ncu launches exit $?"
PROF2="python bench.py --steps 1 --warmup 1 --workload tgv512 --no-e2e --no-cpu-baseline"
$PROF2 > $out/plain2_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$kre -s 40 -c 3 -f -o $out/prof_$tag $PROF2 > $out/ncu_full_$tag.log 2>&1
echo "ncu full exit $?"
ls -la $out
