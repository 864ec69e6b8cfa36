#!/bin/bash
# scripts/gpu_r2k.sh — allocation-free adaptive-dt entry point: alone in a fresh process, then after the whole GPU suite
tag="${1:-r2k}"
out=gpurun_out
mkdir -p $out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $out/pytest_noalloc_alone_$tag.log 2>&1; echo "alone exit $?" | tee -a $out/pytest_noalloc_alone_$tag.log
tail -3 $out/pytest_noalloc_alone_$tag.log
timeout 1200 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_gpu_$tag.log
tail -5 $out/pytest_gpu_$tag.log
timeout 600 python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench512 exit $?"
tail -c 600 $out/bench512_$tag.log
