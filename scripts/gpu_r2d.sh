#!/bin/bash
# scripts/gpu_r2d.sh — 2 GPUs: slab parity tests (big tiles, 3 distributed levels), weak-scaling bench with the parity
# object, strong-scaling channel workload at 1 and 2 GPUs; momentum kernel timing (unrolled register ring)
tag="${1:-r2d}"
out=gpurun_out
mkdir -p $out
nvidia-smi -L > $out/gpus_$tag.txt
nvidia-smi topo -m > $out/topo_$tag.txt 2>&1
timeout 900 python -m pytest tests/test_slab_gloo.py -m gpu -x -q > $out/pytest_slab_$tag.log 2>&1; echo "pytest slab exit $?" >> $out/pytest_slab_$tag.log
tail -4 $out/pytest_slab_$tag.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 2 --no-cpu-baseline > $out/bench512_2gpu_$tag.log 2>&1; echo "bench 2gpu exit $?"
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 2 --no-cpu-baseline --no-parity --workload chan512x256x512 > $out/chan512_2gpu_$tag.log 2>&1; echo "chan512 2gpu exit $?"
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-parity --workload chan512x256x512 > $out/chan512_1gpu_$tag.log 2>&1; echo "chan512 1gpu exit $?"
timeout 600 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu-baseline --no-parity --no-second > $out/bench512_1gpu_$tag.log 2>&1
python - <<PY
import json
for f in ("bench512_2gpu_$tag", "chan512_2gpu_$tag", "chan512_1gpu_$tag", "bench512_1gpu_$tag"):
    try:
        j = json.loads([l for l in open("$out/" + f + ".log") if l.startswith("{")][-1])
        r = j["roofline"]
        print(f, "value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "mom", round(r["momentum"]["ms_per_step"], 2),
              "vcyc ms", round(r["vcycle"]["ms"], 3) if r.get("vcycle") else None, "cycles", j["config"]["vcycles_per_solve"],
              "e2e", round(j["e2e"]["value"], 1) if j.get("e2e") else None, "parity", j.get("parity"))
    except Exception as e:
        print(f, "failed", e)
PY
ls -la $out | tail -8
