#!/bin/bash
# scripts/gpu_r2a.sh — round 2, first call: FP64 pipe peaks, GPU parity tests, closure kernel A/B, headline bench
# with the closure, smoke, ncu capture of the new closure kernel.
tag="${1:-r2a}"
out=gpurun_out
mkdir -p $out
nvidia-smi -L > $out/gpus_$tag.txt
./scripts/_build/fp64_peaks > $out/fp64_peaks_$tag.json 2>&1; echo "fp64_peaks exit $?"; cat $out/fp64_peaks_$tag.json
python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu_$tag.log
tail -5 $out/pytest_gpu_$tag.log
python scripts/nnbench.py 256 256 256 3 > $out/nnbench_$tag.log 2>&1
CFDNN_NN_GENERIC=1 python scripts/nnbench.py 256 256 256 3 >> $out/nnbench_$tag.log 2>&1
python scripts/nnbench.py 512 512 128 3 >> $out/nnbench_$tag.log 2>&1
cat $out/nnbench_$tag.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke_$tag.log 2>&1; echo "smoke exit $?"; tail -4 $out/smoke_$tag.log
python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench512 exit $?"
tail -1 $out/bench512_$tag.log | cut -c1-400
python bench.py --impl reference --steps 2 --warmup 1 > $out/bench_reference_$tag.log 2>&1; echo "bench reference exit $?"
NNP="python scripts/nnbench.py 256 256 64 1"
$NNP > $out/plain_nn_$tag.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_nn_update_mmah -c 1 -f -o $out/prof_nn_$tag $NNP > $out/ncu_nn_$tag.log 2>&1
echo "ncu nn exit $?"
ls -la $out | tail -12
