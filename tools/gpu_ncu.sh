#!/bin/bash
# one ncu --set full capture of k_step mid-episode (launch #SKIP of the bench loop)
tag=${1:-ncu}
skip=${2:-60}
out=gpurun_out/$tag
mkdir -p $out
SHORT="python bench.py --steps $((skip+20)) --warmup 4 --no-cpu-baseline --e2e-steps 3"
timeout 300 $SHORT > $out/plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s $skip -c 1 -o $out/prof_step $SHORT > $out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 $out/ncu_full.log
