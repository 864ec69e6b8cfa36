#!/bin/bash
# scripts/gpu_final1.sh [tag] — final single-GPU check of the round: smoke(), the whole GPU suite, the default bench line
tag="${1:-r2w}"
out=gpurun_out; mkdir -p $out
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $out/smoke_$tag.log 2>&1; echo "smoke exit $?"; tail -2 $out/smoke_$tag.log
timeout 400 python -m pytest tests -m gpu -x -q > $out/pytest_gpu_$tag.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_gpu_$tag.log
tail -4 $out/pytest_gpu_$tag.log
timeout 200 python bench.py --steps 3 --warmup 3 > $out/bench512_$tag.log 2>&1; echo "bench exit $?"
python - <<PY
import json
j = json.loads([l for l in open("$out/bench512_$tag.log") if l.startswith("{")][-1])
r = j["roofline"]
print("value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "frac", round(r["frac"], 3), "nn ms", round(r["nn"]["ms_per_step"], 1),
      "mom", round(r["momentum"]["ms_per_step"], 2), "e2e", round(j["e2e"]["value"], 1), "cpu", j["cpu_baseline"]["value"], "launches", j["gpu_launches"], "parity", j["parity"]["ok"])
PY
