#!/bin/bash
# scripts/gpu_r2b.sh — round 2, second call: table-tanh closure kernel + z-marching momentum kernel:
# parity tests, A/B timings, launch list of the headline step with DRAM bytes, ncu captures of both kernels.
tag="${1:-r2b}"
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu_$tag.log
tail -5 $out/pytest_gpu_$tag.log
python scripts/nnbench.py 256 256 256 3 > $out/nnbench_$tag.log 2>&1
python scripts/nnbench.py 512 512 128 3 >> $out/nnbench_$tag.log 2>&1
cat $out/nnbench_$tag.log
python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench512 exit $?"
tail -1 $out/bench512_$tag.log | cut -c1-200
CFDNN_MOMENTUM_Z=0 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu-baseline --no-parity --no-second > $out/bench512_oldmom_$tag.log 2>&1
CFDNN_MOMENTUM_EXACT=1 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu-baseline --no-parity --no-second > $out/bench512_exactmom_$tag.log 2>&1
python - <<PY
import json
for f in ("bench512_$tag", "bench512_oldmom_$tag", "bench512_exactmom_$tag"):
    try:
        j = json.loads([l for l in open("$out/" + f + ".log") if l.startswith("{")][-1])
        print(f, "ms/step", round(j["ms_per_step"], 2), "momentum ms/step", round(j["roofline"]["momentum"]["ms_per_step"], 3),
              "nn ms", round(j["roofline"]["nn"]["ms_per_step"], 2), "vcyc ms", round(j["roofline"]["vcycle"]["ms"], 3))
    except Exception as e:
        print(f, "failed", e)
PY
PROF="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-parity --no-second"
$PROF > $out/plain_$tag.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 12000 --csv --log-file $out/launches512_$tag.csv $PROF > $out/ncu_launches_$tag.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k regex:k_momentum_z -s 3 -c 1 -f -o $out/prof_mom_$tag $PROF > $out/ncu_mom_$tag.log 2>&1
echo "ncu momentum exit $?"
NNP="python scripts/nnbench.py 256 256 64 1"
$NNP > $out/plain_nn_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_nn_update_mmah -c 1 -f -o $out/prof_nn_$tag $NNP > $out/ncu_nn_$tag.log 2>&1
echo "ncu nn exit $?"
gzip -f $out/launches512_$tag.csv
ls -la $out | tail -14
