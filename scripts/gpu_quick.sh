#!/bin/bash
# scripts/gpu_quick.sh tag [ENV=VAL for the A/B leg] — GPU test suite, headline bench, and the same bench with one switch flipped
tag="${1:-q}"; ab="${2:-}"; out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $? $(tail -1 $out/pytest_gpu_$tag.log)"
timeout 600 python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench exit $?"; tail -1 $out/bench512_$tag.log | cut -c1-160
if [ -n "$ab" ]; then
  env $ab timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $out/bench512_${tag}_ab.log 2>&1; echo "bench ($ab) exit $?"; tail -1 $out/bench512_${tag}_ab.log | cut -c1-160
fi
