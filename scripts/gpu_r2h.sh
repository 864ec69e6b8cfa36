#!/bin/bash
# scripts/gpu_r2h.sh — y-wall patch smoother on channel workloads, config 3 (stretched) bench, adaptive-dt bench with
# launch list, full GPU test suite
tag="${1:-r2h}"
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu_$tag.log
tail -4 $out/pytest_gpu_$tag.log
for m in 2 3; do
  for n in 256; do
    CFDNN_MG_TBO_MODE=$m timeout 120 python scripts/mgbench.py $n 4 3 PPNNPP 2>&1 | tail -1 | sed "s/^/mode $m /" | tee -a $out/mgbench_chan_$tag.log
  done
done
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-parity --workload chan512x256x512 > $out/chan512_1gpu_$tag.log 2>&1; echo "chan512 exit $?"
CFDNN_MG_TBO_MODE=2 timeout 600 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu-baseline --no-parity --workload chan512x256x512 > $out/chan512_1gpu_general_$tag.log 2>&1
timeout 600 python bench.py --steps 2 --warmup 1 --no-parity --workload chan192x96x192 > $out/chan192_$tag.log 2>&1; echo "chan192 (config 3) exit $?"
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 --workload chan192x96x192 > $out/chan192_reference_$tag.log 2>&1; echo "chan192 reference exit $?"
timeout 600 python bench.py --steps 3 --warmup 2 --adaptive-dt --no-cpu-baseline --no-parity --no-second > $out/bench512_adaptive_$tag.log 2>&1; echo "adaptive exit $?"
python - <<PY
import json, glob
for f in ("chan512_1gpu_$tag", "chan512_1gpu_general_$tag", "chan192_$tag", "chan192_reference_$tag", "bench512_adaptive_$tag"):
    try:
        j = json.loads([l for l in open("$out/" + f + ".log") if l.startswith("{")][-1])
        r = j.get("roofline") or {}
        print(f, "value", round(j["value"], 3), "ms/step", round(j["ms_per_step"], 2), "vcyc ms", round(r["vcycle"]["ms"], 3) if r.get("vcycle") else None,
              "cycles", j["config"].get("vcycles_per_solve"), "cpu", (j.get("cpu_baseline") or {}).get("value"), "dt", j["config"].get("dt_last"))
    except Exception as e:
        print(f, "failed", e)
PY
PROF="python bench.py --steps 1 --warmup 1 --adaptive-dt --no-cpu-baseline --no-e2e --no-parity --no-second"
$PROF > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 12000 --csv --log-file $out/launches512_adaptive_$tag.csv $PROF > $out/ncu_launches_$tag.log 2>&1
echo "ncu launches exit $?"
gzip -f $out/launches512_adaptive_$tag.csv
ls -la $out | tail -8
