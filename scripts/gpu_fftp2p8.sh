#!/bin/bash
# scripts/gpu_fftp2p8.sh N [tag] — peer-to-peer FFT transposes at N GPUs: slab parity (both routes) + config 5
n="${1:-8}"; tag="${2:-r2v}"
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_slab_gloo.py -m gpu -x -q -k "fft and $n" > $out/pytest_slabfft_p2p_${n}gpu_$tag.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_slabfft_p2p_${n}gpu_$tag.log
tail -3 $out/pytest_slabfft_p2p_${n}gpu_$tag.log
NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
  bench.py --gpus $n --steps 3 --warmup 2 --workload chan1024x512x1024 --poisson fft --no-cpu-baseline > $out/bench_chan1024x512x1024_fft_p2p_${n}gpu_$tag.log 2>&1
echo "bench exit $?"
python - <<PY
import json
j = json.loads([l for l in open("$out/bench_chan1024x512x1024_fft_p2p_${n}gpu_$tag.log") if l.startswith("{")][-1])
r = j["roofline"]
print("x$n value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "poisson", round(r["poisson"]["ms_per_step"], 2), r["poisson"].get("transposes"),
      "mom", round(r["momentum"]["ms_per_step"], 2), "parity", (j.get("parity") or {}).get("ok"), "e2e", round((j.get("e2e") or {}).get("value") or 0, 1))
PY
