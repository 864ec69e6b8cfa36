#!/bin/bash
# scripts/gpu_suite.sh [tag] — the whole GPU suite (+ the allocation-free entry point alone in a fresh process)
tag="${1:-r2r}"
out=gpurun_out
mkdir -p $out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $out/pytest_noalloc_alone_$tag.log 2>&1; echo "alone exit $?" | tee -a $out/pytest_noalloc_alone_$tag.log
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_gpu_$tag.log
tail -16 $out/pytest_gpu_$tag.log
