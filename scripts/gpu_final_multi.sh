#!/bin/bash
# scripts/gpu_final_multi.sh N [tag] [full] — one gpurun --gpus N call: slab parity tests over NCCL at N GPUs (multigrid big
# tiles + FFT transposes) and BASELINE config 5 (1024x512x1024 channel, strong scaling) with the FFT solver and the multigrid.
n="${1:-8}"; tag="${2:-r2s}"
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_slab_gloo.py -m gpu -x -q -k "$n" > $out/pytest_slab_${n}gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_slab_${n}gpu_$tag.log
tail -4 $out/pytest_slab_${n}gpu_$tag.log
run() {  # workload, poisson
  NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $n --steps 3 --warmup 2 --workload $1 --poisson $2 --no-cpu-baseline > $out/bench_$1_$2_${n}gpu_$tag.log 2>&1
  echo "bench $1 $2 x$n exit $?"
  python - <<PY
import json
try:
    j = json.loads([l for l in open("$out/bench_$1_$2_${n}gpu_$tag.log") if l.startswith("{")][-1])
    r = j["roofline"]
    print("$1 $2 x$n value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "poisson", round(r["poisson"]["ms_per_step"], 2),
          "mom", round(r["momentum"]["ms_per_step"], 2), "vcyc", (r.get("vcycle") or {}).get("ms"), "cycles", j["config"]["vcycles_per_solve"],
          "parity", (j.get("parity") or {}).get("ok"), "e2e", round((j.get("e2e") or {}).get("value") or 0, 1))
except Exception as e:
    print("parse failed", e)
PY
}
run chan1024x512x1024 fft
if [ "${3:-}" = "full" ]; then run chan1024x512x1024 mg; fi
