#!/bin/bash
# scripts/gpu_multi.sh N [tag] — one gpurun --gpus N call: slab parity tests over NCCL + weak-scaling bench at N GPUs.
n="${1:-2}"; tag="${2:-r1}"
out=gpurun_out; mkdir -p $out
nvidia-smi topo -m > $out/topo_${n}gpu_$tag.txt 2>&1
timeout 900 python -m pytest tests/test_slab_gloo.py -m gpu -x -q > $out/pytest_slab_${n}gpu_$tag.log 2>&1; echo "pytest exit $?" >> $out/pytest_slab_${n}gpu_$tag.log
tail -4 $out/pytest_slab_${n}gpu_$tag.log
for wl in tgv512 tgv256; do
  NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $n --steps 3 --warmup 3 --workload $wl --no-cpu-baseline > $out/bench_${wl}_${n}gpu_$tag.log 2>&1
  echo "bench $wl x$n exit $?"; tail -1 $out/bench_${wl}_${n}gpu_$tag.log | cut -c1-700
done
