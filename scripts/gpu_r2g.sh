#!/bin/bash
# scripts/gpu_r2g.sh — TMA-fed smoother (k_sweep_tbx, modes 4/5) vs k_sweep_tbq (mode 3): variant equality, timing, ncu;
# momentum kernel profile (unrolled ring)
tag="${1:-r2g}"
out=gpurun_out
mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "variants" > $out/pytest_variants_$tag.log 2>&1; echo "variants exit $?"
tail -3 $out/pytest_variants_$tag.log
for m in 3 4 5; do
  for n in 512 256 128; do
    CFDNN_MG_TBO_MODE=$m timeout 120 python scripts/mgbench.py $n 4 3 2>&1 | tail -1 | sed "s/^/mode $m /" | tee -a $out/mgbench_$tag.log
  done
done
for m in 4 5; do
  CFDNN_MG_TBO_MODE=$m timeout 600 python -m pytest tests -m gpu -x -q -k "mg_ or tgv or slab or step or golden" > $out/pytest_gpu_m${m}_$tag.log 2>&1; echo "mode $m pytest exit $? $(tail -1 $out/pytest_gpu_m${m}_$tag.log)"
  CFDNN_MG_TBO_MODE=$m timeout 600 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu-baseline --no-second > $out/bench512_m${m}_$tag.log 2>&1
done
python - <<PY
import json
for f in ("bench512_m4_$tag", "bench512_m5_$tag"):
    try:
        j = json.loads([l for l in open("$out/" + f + ".log") if l.startswith("{")][-1])
        r = j["roofline"]
        print(f, "value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "smoother launch ms", round(r["launch_ms"], 4), "frac", round(r["frac"], 3),
              "vcyc ms", round(r["vcycle"]["ms"], 3), "parity", j["parity"]["ok"])
    except Exception as e:
        print(f, "failed", e)
PY
MGP="python scripts/mgbench.py 512 2 1"
CFDNN_MG_TBO_MODE=4 $MGP > $out/plain_mg_$tag.log 2>&1 &&
CFDNN_MG_TBO_MODE=4 ncu --set full --clock-control none --import-source on -k regex:k_sweep_tbx -s 2 -c 1 -f -o $out/prof_tbx_$tag $MGP > $out/ncu_tbx_$tag.log 2>&1
echo "ncu tbx exit $?"
PROF="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-parity --no-second"
ncu --set full --clock-control none --import-source on -k regex:k_momentum_z -s 3 -c 1 -f -o $out/prof_mom_$tag $PROF > $out/ncu_mom_$tag.log 2>&1
echo "ncu momentum exit $?"
ls -la $out | tail -8
