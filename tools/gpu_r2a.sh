#!/bin/bash
# round-2 call A: full GPU test suite (incl. no-resync scale tests), bench both arms, E sweep, cfg B/C, ncu evidence
tag=${1:-r2_a}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $out/clocks.csv &
SMI=$!
timeout 1500 python -m pytest tests -m gpu -x -q -s > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest_gpu.log
grep -E "no-resync|passed|failed|Error" $out/pytest_gpu.log | tail -15
timeout 600 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"
tail -c 3000 $out/bench.json
timeout 400 python bench.py --impl reference --steps 20 --warmup 5 > $out/bench_ref.json 2>> $out/bench.err; echo "ref rc=$?"
cat $out/bench_ref.json
for E in 1024 2048 8192 16384 32768; do
  timeout 300 python bench.py --envs-per-gpu $E --steps 360 --warmup 40 --no-cpu-baseline --no-other-configs --e2e-steps 3 > $out/sweep_$E.json 2>> $out/bench.err
  echo "sweep $E rc=$?"
done
kill $SMI
timeout 300 python tools/phase_clocks.py > $out/phase.log 2>&1
timeout 300 python tools/plan_phase_clocks.py > $out/plan_phase.log 2>&1; cat $out/plan_phase.log
timeout 300 python tools/step_profile.py > $out/step_profile.txt 2>&1
SHORT="python bench.py --steps 12 --warmup 4 --no-cpu-baseline --no-other-configs --e2e-steps 3"
timeout 300 $SHORT > $out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $out/launches.csv $SHORT > $out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_plan$ -s 30 -c 1 -o $out/prof_plan $SHORT > $out/ncu_full_plan.log 2>&1
echo "ncu full k_plan rc=$?"
B100="python bench.py --scenario 100v100 --steps 80 --warmup 4 --no-cpu-baseline --no-other-configs --e2e-steps 3"
timeout 300 $B100 > $out/plain100.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 60 -c 1 -o $out/prof_step100 $B100 > $out/ncu_full_step100.log 2>&1
echo "ncu full k_step 100v100 rc=$?"
LONG="python bench.py --steps 80 --warmup 4 --no-cpu-baseline --no-other-configs --e2e-steps 3"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 60 -c 1 -o $out/prof_step $LONG > $out/ncu_full.log 2>&1
echo "ncu full k_step rc=$?"
for k in observe step; do
  R="python tools/dram_range.py --kernel $k --launches 64"
  timeout 300 $R > $out/range_${k}_plain.log 2>&1 &&
  timeout 600 ncu --replay-mode range --profile-from-start off --cache-control none --clock-control none \
     --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv --log-file $out/range_$k.csv $R > $out/range_${k}_ncu.log 2>&1
  echo "range $k rc=$?"; cat $out/range_${k}_plain.log | tail -1; tail -5 $out/range_$k.csv
done
ls -la $out
