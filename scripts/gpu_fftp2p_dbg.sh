#!/bin/bash
# scripts/gpu_fftp2p_dbg.sh N — diagnose the peer-to-peer transposes: IPC error strings, Python stack on a fault
n="${1:-2}"; out=gpurun_out; mkdir -p $out
for mode in 1 0; do
  env CFDNN_FFT_P2P=$mode CFDNN_FFT_P2P_DEBUG=1 PYTHONFAULTHANDLER=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n \
    --master-addr 127.0.0.1 --master-port $((29540 + mode)) bench.py --gpus $n --steps 2 --warmup 1 --workload chan512x256x512 --poisson fft \
    --no-cpu-baseline --no-e2e --no-parity > $out/p2p_dbg_$mode.log 2>&1
  echo "mode $mode exit $?"
  grep -E "cfdnn fft p2p|Fatal|File \"|value|Segmentation|line [0-9]+ in" $out/p2p_dbg_$mode.log | cut -c1-260 | head -30
done
