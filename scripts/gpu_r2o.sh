#!/bin/bash
# scripts/gpu_r2o.sh — FFT Poisson with batched loads; launch list of a channel step with the FFT solver
tag="${1:-r2o}"
out=gpurun_out
mkdir -p $out
{
python scripts/fftbench.py 512 256 512
python scripts/fftbench.py 256 128 256
python scripts/fftbench.py 192 96 192
} > $out/fftbench_$tag.log 2>&1
cat $out/fftbench_$tag.log
PROF="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-parity --workload chan512x256x512 --poisson fft"
$PROF > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $out/launches_chanfft_$tag.csv $PROF > $out/ncu_list_$tag.log 2>&1
echo "ncu list exit $?"
python scripts/agg_launches.py $out/launches_chanfft_$tag.csv 2>&1 | head -40
