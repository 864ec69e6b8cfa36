#!/bin/bash
# scripts/gpu_r2l.sh — FFT Poisson solver: GPU tests, channel bench with mg vs fft, launch list + ncu of the FFT kernels
tag="${1:-r2l}"
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fft" > $out/pytest_fft_$tag.log 2>&1; echo "pytest fft exit $?" | tee -a $out/pytest_fft_$tag.log
tail -5 $out/pytest_fft_$tag.log
for wl in chan512x256x512 chan256x128x256; do
  for ps in mg fft; do
    timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-parity --no-e2e --workload $wl --poisson $ps > $out/${wl}_${ps}_$tag.log 2>&1; echo "$wl $ps exit $?"
  done
done
python - <<PY
import json
for wl in ("chan512x256x512", "chan256x128x256"):
    for ps in ("mg", "fft"):
        f = f"$out/{wl}_{ps}_$tag.log"
        try:
            j = json.loads([l for l in open(f) if l.startswith("{")][-1])
            r = j["roofline"]
            print(wl, ps, "value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "poisson", r.get("poisson"), "mom", round(r["momentum"]["ms_per_step"], 2))
        except Exception as e:
            print(f, "failed", e)
PY
PROF="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-parity --workload chan512x256x512 --poisson fft"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_fft" -c 40 --csv --log-file $out/launches_fft_$tag.csv $PROF > $out/ncu_fft_list_$tag.log 2>&1
echo "ncu list exit $?"
grep "k_fft" $out/launches_fft_$tag.csv | awk -F'","' '{print $5, $(NF-2), $(NF)}' | tail -24
ncu --set full --clock-control none --import-source on -k regex:"k_fft_x_fwd|k_fft_z|k_fft_thomas" -s 8 -c 3 -o $out/prof_fft_$tag -f $PROF > $out/ncu_fft_$tag.log 2>&1
echo "ncu full exit $?"
ls -la $out | tail -8
