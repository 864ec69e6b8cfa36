#!/bin/bash
# scripts/gpu_r2c.sh — 12-instruction tanh + software-pipelined momentum kernel: tests, timings, ncu of the momentum kernel
tag="${1:-r2c}"
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu_$tag.log
tail -4 $out/pytest_gpu_$tag.log
python scripts/nnbench.py 256 256 256 3 > $out/nnbench_$tag.log 2>&1
cat $out/nnbench_$tag.log
python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench512 exit $?"
CFDNN_MOMENTUM_Z=0 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu-baseline --no-parity --no-second > $out/bench512_oldmom_$tag.log 2>&1
python - <<PY
import json
for f in ("bench512_$tag", "bench512_oldmom_$tag"):
    try:
        j = json.loads([l for l in open("$out/" + f + ".log") if l.startswith("{")][-1])
        print(f, "ms/step", round(j["ms_per_step"], 2), "momentum ms/step", round(j["roofline"]["momentum"]["ms_per_step"], 3),
              "nn ms", round(j["roofline"]["nn"]["ms_per_step"], 2), "vcyc ms", round(j["roofline"]["vcycle"]["ms"], 3))
    except Exception as e:
        print(f, "failed", e)
PY
PROF="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-parity --no-second"
$PROF > $out/plain_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_momentum_z -s 3 -c 1 -f -o $out/prof_mom_$tag $PROF > $out/ncu_mom_$tag.log 2>&1
echo "ncu momentum exit $?"
ls -la $out | tail -8
