#!/bin/bash
# scripts/gpu_r2n.sh — FFT Poisson: radix-8 on/off timings, tests, ncu
tag="${1:-r2n}"
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fft" > $out/pytest_fft_$tag.log 2>&1; echo "pytest fft exit $?" | tee -a $out/pytest_fft_$tag.log
tail -3 $out/pytest_fft_$tag.log
{
python scripts/fftbench.py 512 256 512
CFDNN_FFT_RADIX8=0 python scripts/fftbench.py 512 256 512
python scripts/fftbench.py 256 128 256
CFDNN_FFT_RADIX8=0 python scripts/fftbench.py 256 128 256
python scripts/fftbench.py 192 96 192
python scripts/fftbench.py 1024 512 1024 4
} > $out/fftbench_$tag.log 2>&1
cat $out/fftbench_$tag.log
PROF="python scripts/fftbench.py 512 256 512 2"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_fft" -s 24 -c 6 --csv --log-file $out/launches_fft_$tag.csv $PROF > $out/ncu_fft_list_$tag.log 2>&1
grep "k_fft" $out/launches_fft_$tag.csv | grep gpu__time | awk -F'","' '{print $5, $(NF)}'
ncu --set full --clock-control none --import-source on -k regex:"k_fft_x_fwd|k_fft_z|k_fft_thomas|k_fft_x_inv" -s 16 -c 5 -o $out/prof_fft_$tag -f $PROF > $out/ncu_fft_$tag.log 2>&1
echo "ncu full exit $?"
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-parity --no-e2e --workload chan512x256x512 --poisson fft > $out/chan512_fft_$tag.log 2>&1; echo "chan512 fft exit $?"
python - <<PY
import json
j = json.loads([l for l in open("$out/chan512_fft_$tag.log") if l.startswith("{")][-1])
print("chan512 fft", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), j["roofline"]["poisson"]["ms_per_step"], j["roofline"]["momentum"]["ms_per_step"])
PY
