#!/bin/bash
# scripts/gpu_r2e.sh — 8 GPUs: slab parity tests at 2/4/8 ranks, weak scaling (tgv512 per GPU) at 4 and 8,
# strong scaling of BASELINE config 5 (1024x512x1024 channel) at 2/4/8.
tag="${1:-r2e}"
out=gpurun_out
mkdir -p $out
nvidia-smi -L > $out/gpus_$tag.txt
nvidia-smi topo -m > $out/topo_$tag.txt 2>&1
timeout 1500 python -m pytest tests/test_slab_gloo.py -m gpu -x -q > $out/pytest_slab_$tag.log 2>&1; echo "pytest slab exit $?" >> $out/pytest_slab_$tag.log
tail -4 $out/pytest_slab_$tag.log
run() {  # n workload extra...
  n=$1; wl=$2; shift 2
  devs=$(seq -s, 0 $((n-1)))
  CUDA_VISIBLE_DEVICES=$devs timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) \
      bench.py --gpus $n --steps 3 --warmup 2 --no-cpu-baseline --workload $wl "$@" > $out/${wl}_${n}gpu_$tag.log 2>&1
  echo "$wl x$n exit $?"
}
run 8 tgv512
run 4 tgv512
run 8 chan1024x512x1024 --no-parity
run 4 chan1024x512x1024 --no-parity
run 2 chan1024x512x1024 --no-parity
python - <<PY
import json, glob
for f in sorted(glob.glob("$out/*gpu_$tag.log")):
    try:
        j = json.loads([l for l in open(f) if l.startswith("{")][-1])
        r = j["roofline"]
        print(f.split("/")[-1], "value", round(j["value"], 1), "ms/step", round(j["ms_per_step"], 2), "vcyc ms", round(r["vcycle"]["ms"], 3),
              "cycles", j["config"]["vcycles_per_solve"], "e2e", round(j["e2e"]["value"], 1) if j.get("e2e") else None,
              "e2e ms", {k: round(v, 1) for k, v in j["e2e"]["ms_per_step"].items()} if j.get("e2e") else None,
              "numa", j["e2e"].get("host_numa") if j.get("e2e") else None, "parity", (j.get("parity") or {}).get("ok"), (j.get("parity") or {}).get("max_rel_diff_vs_single_domain"))
    except Exception as e:
        print(f, "failed", e)
PY
ls -la $out | tail -12
