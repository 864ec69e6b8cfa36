#!/bin/bash
# scripts/gpu_r2i.sh — implicit y-diffusion, wall-row patch smoother, row-strided CFL maxima: GPU tests + timings
tag="${1:-r2i}"
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu_$tag.log
tail -4 $out/pytest_gpu_$tag.log
timeout 600 python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench512 exit $?"
timeout 600 python bench.py --steps 3 --warmup 2 --adaptive-dt --no-cpu-baseline --no-parity --no-second --no-e2e > $out/bench512_adaptive_$tag.log 2>&1; echo "adaptive exit $?"
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-parity --no-e2e --workload chan512x256x512 --implicit-y > $out/chan512_implicit_$tag.log 2>&1; echo "chan512 implicit exit $?"
python - <<PY
import json
for f in ("bench512_$tag", "bench512_adaptive_$tag", "chan512_implicit_$tag"):
    try:
        j = json.loads([l for l in open("$out/" + f + ".log") if l.startswith("{")][-1])
        r = j["roofline"]
        print(f, "value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "smoother ms", round(r["launch_ms"], 4), "mom", round(r["momentum"]["ms_per_step"], 2),
              "vcyc", round(r["vcycle"]["ms"], 3), "cycles", j["config"]["vcycles_per_solve"], "implicit", r.get("implicit_y"), "e2e", (j.get("e2e") or {}).get("value"))
    except Exception as e:
        print(f, "failed", e)
PY
PROF="python bench.py --steps 1 --warmup 1 --adaptive-dt --no-cpu-baseline --no-e2e --no-parity --no-second"
$PROF > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_cfl_maxima|k_final_max_rows|k_thomas" -c 40 --csv --log-file $out/launches_cfl_$tag.csv $PROF > $out/ncu_cfl_$tag.log 2>&1
echo "ncu cfl exit $?"
grep -E "k_cfl_maxima|k_final_max_rows" $out/launches_cfl_$tag.csv | awk -F'","' '{print $5, $(NF)}' | head -8
ls -la $out | tail -6
