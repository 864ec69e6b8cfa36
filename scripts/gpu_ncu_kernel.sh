#!/bin/bash
# scripts/gpu_ncu_kernel.sh regex tag [N] — ncu --set full capture of one kernel inside a 2-cycle multigrid solve (mgbench)
re="$1"; tag="${2:-k}"; n="${3:-512}"; out=gpurun_out; mkdir -p $out
python scripts/mgbench.py $n 2 1 > $out/mgbench_plain_$tag.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:$re -s 1 -c 1 -f -o $out/prof_$tag python scripts/mgbench.py $n 2 1 > $out/ncu_$tag.log 2>&1
echo "ncu exit $?"; tail -2 $out/ncu_$tag.log
