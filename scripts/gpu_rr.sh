#!/bin/bash
# scripts/gpu_rr.sh tag — fused residual+restriction: parity tests, V-cycle A/B (CFDNN_MG_RR=0/1), bench, ncu of the fused kernel
tag="${1:-rr}"; out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $out/pytest_$tag.log 2>&1; echo "pytest exit $? $(tail -1 $out/pytest_$tag.log)"
for rr in 0 1; do
  for n in 512 256; do
    CFDNN_MG_RR=$rr timeout 300 python scripts/mgbench.py $n 4 3 2>&1 | tail -1 | sed "s/^/rr $rr /" | tee -a $out/mgbench_$tag.log
  done
done
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $out/bench512_$tag.log 2>&1; echo "bench exit $?"; tail -1 $out/bench512_$tag.log | cut -c1-200
ncu --set full --clock-control none --import-source on -k regex:k_residual_restrict -c 1 -f -o $out/prof_$tag python scripts/mgbench.py 512 2 1 > $out/ncu_$tag.log 2>&1; echo "ncu exit $?"
ncu --set full --clock-control none --import-source on -k regex:k_momentum -c 1 -f -o $out/prof_mom_$tag python bench.py --steps 1 --warmup 0 --no-cpu-baseline --no-e2e > $out/ncu_mom_$tag.log 2>&1; echo "ncu mom exit $?"
