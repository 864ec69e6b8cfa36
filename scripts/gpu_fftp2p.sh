#!/bin/bash
# scripts/gpu_fftp2p.sh N [tag] — FFT Poisson on slabs: peer-to-peer transposes (CUDA IPC stores from the transform kernels)
# against NCCL all-to-alls: slab parity test, config 5 bench both ways.
n="${1:-2}"; tag="${2:-r2t}"
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_slab_gloo.py -m gpu -x -q -k "fft and $n" > $out/pytest_slabfft_p2p_${n}gpu_$tag.log 2>&1; echo "pytest p2p exit $?" | tee -a $out/pytest_slabfft_p2p_${n}gpu_$tag.log
tail -3 $out/pytest_slabfft_p2p_${n}gpu_$tag.log
run() {  # tag suffix, env
  env $2 NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $n --steps 3 --warmup 2 --workload chan1024x512x1024 --poisson fft --no-cpu-baseline --no-e2e > $out/bench_chan1024_fft_$1_${n}gpu_$tag.log 2>&1
  echo "bench $1 x$n exit $?"
  python - <<PY
import json
try:
    j = json.loads([l for l in open("$out/bench_chan1024_fft_$1_${n}gpu_$tag.log") if l.startswith("{")][-1])
    r = j["roofline"]
    print("$1 x$n value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "poisson", round(r["poisson"]["ms_per_step"], 2),
          r["poisson"].get("transposes"), "parity", (j.get("parity") or {}).get("ok"))
except Exception as e:
    print("parse failed", e)
PY
}
run p2p CFDNN_FFT_P2P=1
run nccl CFDNN_FFT_P2P=0
