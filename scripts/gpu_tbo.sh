#!/bin/bash
# scripts/gpu_tbo.sh tag "modes" "pfds" — blocked smoother variants (CFDNN_MG_TBO_MODE x CFDNN_MG_TBO_PFD): tests + V-cycle timing
tag="${1:-tbo}"; modes="${2:-0 2}"; pfds="${3:-0}"; out=gpurun_out; mkdir -p $out
for m in $modes; do
  for p in $pfds; do
    export CFDNN_MG_TBO_MODE=$m CFDNN_MG_TBO_PFD=$p
    timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "variants or mg_solve or tgv64 or rerun" > $out/pytest_${tag}_m${m}_p$p.log 2>&1; echo "mode $m pfd $p pytest exit $? $(tail -1 $out/pytest_${tag}_m${m}_p$p.log)"
    for n in 512 256; do
      timeout 300 python scripts/mgbench.py $n 4 3 2>&1 | tail -1 | sed "s/^/mode $m pfd $p /" | tee -a $out/mgbench_$tag.log
    done
  done
done
