#!/bin/bash
# scripts/gpu_r2f.sh — 2 x 2 patch smoother (k_sweep_tbq) vs k_sweep_tbo: parity tests, V-cycle timing, headline bench, ncu
tag="${1:-r2f}"
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu_$tag.log
tail -4 $out/pytest_gpu_$tag.log
for m in 2 3; do
  for n in 512 256 128; do
    CFDNN_MG_TBO_MODE=$m timeout 300 python scripts/mgbench.py $n 4 3 2>&1 | tail -1 | sed "s/^/mode $m /" | tee -a $out/mgbench_$tag.log
  done
done
timeout 600 python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench512 exit $?"
CFDNN_MG_TBO_MODE=2 timeout 600 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu-baseline --no-parity --no-second > $out/bench512_tbo_$tag.log 2>&1
python - <<PY
import json
for f in ("bench512_$tag", "bench512_tbo_$tag"):
    try:
        j = json.loads([l for l in open("$out/" + f + ".log") if l.startswith("{")][-1])
        r = j["roofline"]
        print(f, "value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "smoother launch ms", round(r["launch_ms"], 4), "frac", round(r["frac"], 3),
              "momentum", round(r["momentum"]["ms_per_step"], 2), "nn", round(r["nn"]["ms_per_step"], 2), "vcyc ms", round(r["vcycle"]["ms"], 3))
    except Exception as e:
        print(f, "failed", e)
PY
MGP="python scripts/mgbench.py 512 2 1"
$MGP > $out/plain_mg_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_sweep_tbq -s 2 -c 1 -f -o $out/prof_tbq_$tag $MGP > $out/ncu_tbq_$tag.log 2>&1
echo "ncu tbq exit $?"
ls -la $out | tail -8
