#!/bin/bash
# scripts/gpu_r2j.sh — LES SGS closures + xz-template momentum: GPU tests, bench with each closure, ncu of k_les_sgs
tag="${1:-r2j}"
out=gpurun_out
mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu_$tag.log
tail -4 $out/pytest_gpu_$tag.log
timeout 600 python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench512 exit $?"
for c in smagorinsky wale vreman sigma; do
  timeout 600 python bench.py --steps 3 --warmup 2 --closure $c --no-cpu-baseline --no-parity --no-e2e > $out/bench512_${c}_$tag.log 2>&1; echo "$c exit $?"
done
python - <<PY
import json
for f in ("bench512_$tag", "bench512_smagorinsky_$tag", "bench512_wale_$tag", "bench512_vreman_$tag", "bench512_sigma_$tag"):
    try:
        j = json.loads([l for l in open("$out/" + f + ".log") if l.startswith("{")][-1])
        r = j["roofline"]
        print(f, "value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "smoother ms", round(r["launch_ms"], 4), "mom", round(r["momentum"]["ms_per_step"], 2),
              "vcyc", round(r["vcycle"]["ms"], 3), "cycles", j["config"]["vcycles_per_solve"], "les", r.get("les"), "e2e", (j.get("e2e") or {}).get("value"))
    except Exception as e:
        print(f, "failed", e)
PY
PROF="python bench.py --steps 1 --warmup 1 --closure wale --no-cpu-baseline --no-e2e --no-parity --no-second"
ncu --set full --clock-control none --import-source on -k regex:"k_les_sgs" -c 1 -o $out/prof_les_$tag -f $PROF > $out/ncu_les_$tag.log 2>&1
echo "ncu les exit $?"
ls -la $out | tail -8
