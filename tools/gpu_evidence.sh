#!/bin/bash
# round-end evidence for the current build: launch list of the bench loop + ncu --set full of k_step, k_observe, k_plan,
# k_policy_encode, k_conc_attend (each after the same command has run clean without ncu), phase clocks, range DRAM counters
tag=${1:-ev}
out=gpurun_out/$tag
mkdir -p $out
SHORT="python bench.py --steps 12 --warmup 4 --no-cpu-baseline --no-other-configs --e2e-steps 3"
timeout 300 $SHORT > $out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $out/launches.csv $SHORT > $out/ncu_list.log 2>&1
echo "ncu list rc=$?"
LONG="python bench.py --steps 80 --warmup 4 --no-cpu-baseline --no-other-configs --e2e-steps 3"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 260 -c 1 -o $out/prof_step $LONG > $out/ncu_full_step.log 2>&1
echo "ncu full k_step rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_observe -s 20 -c 1 -o $out/prof_observe $LONG > $out/ncu_full_observe.log 2>&1
echo "ncu full k_observe rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_plan$ -s 200 -c 1 -o $out/prof_plan $LONG > $out/ncu_full_plan.log 2>&1
echo "ncu full k_plan rc=$?"
POL="python tools/policy_time.py"
timeout 300 $POL > $out/policy_time.txt 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_policy_encode -s 3 -c 1 -o $out/prof_encode $POL > $out/ncu_full_encode.log 2>&1
echo "ncu full k_policy_encode rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_conc_attend -s 6 -c 1 -o $out/prof_attend $POL > $out/ncu_full_attend.log 2>&1
echo "ncu full k_conc_attend rc=$?"
timeout 300 python tools/phase_clocks.py > $out/phase.log 2>&1
timeout 300 python tools/plan_phase_clocks.py > $out/plan_phase.log 2>&1
timeout 300 python tools/step_profile.py > $out/step_profile.txt 2>&1
for k in observe step; do
  R="python tools/dram_range.py --kernel $k --launches 64"
  timeout 300 $R > $out/range_${k}_plain.log 2>&1 &&
  timeout 600 ncu --replay-mode range --cache-control none --clock-control none \
     --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv --log-file $out/range_$k.csv $R > $out/range_${k}_ncu.log 2>&1
  echo "range $k rc=$?"
done
ls -la $out | head -40
