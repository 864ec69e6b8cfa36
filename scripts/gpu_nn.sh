#!/bin/bash
# scripts/gpu_nn.sh — one gpurun call: NN closure parity tests, nnbench (tensor-core and scalar kernels),
# ncu full capture of the tensor-core kernel, ncu full capture of the level-0 blocked smoother.
tag="${1:-r1d}"
out=gpurun_out
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_nn.py -x -q > $out/pytest_nn_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_nn_$tag.log
tail -5 $out/pytest_nn_$tag.log
timeout 300 python scripts/nnbench.py 256 256 256 3 > $out/nnbench_$tag.log 2>&1
CFDNN_NN_SIMT=1 timeout 300 python scripts/nnbench.py 256 256 256 2 >> $out/nnbench_$tag.log 2>&1
timeout 300 python scripts/nnbench.py 512 512 128 2 >> $out/nnbench_$tag.log 2>&1
cat $out/nnbench_$tag.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_nn_update_mma -c 1 -f -o $out/prof_nn_$tag python scripts/nnbench.py 256 256 64 1 > $out/ncu_nn_$tag.log 2>&1
echo "ncu nn exit $?"
PROF2="python bench.py --steps 1 --warmup 1 --workload tgv512 --no-e2e --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_sweep_tb -s 4 -c 2 -f -o $out/prof_tb512_$tag $PROF2 > $out/ncu_tb512_$tag.log 2>&1
echo "ncu tb exit $?"
ls -la $out
