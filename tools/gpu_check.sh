#!/bin/bash
# One gpurun call: GPU parity tests, bench line, phase clocks, ncu launch list + one --set full capture of k_step.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_check.sh <tag>'
tag=${1:-run}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $out/clocks.csv &
SMI=$!
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest_gpu.log
tail -3 $out/pytest_gpu.log
timeout 600 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"
cat $out/bench.json
timeout 300 python bench.py --impl reference --steps 60 --warmup 5 > $out/bench_ref.json 2>> $out/bench.err; echo "ref rc=$?"
cat $out/bench_ref.json
timeout 300 python tools/phase_clocks.py > $out/phase.log 2>&1; cat $out/phase.log
kill $SMI
SHORT="python bench.py --steps 12 --warmup 4 --no-cpu-baseline --e2e-steps 3"
timeout 300 $SHORT > $out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $out/launches.csv $SHORT > $out/ncu_list.log 2>&1
echo "ncu list rc=$?"
LONG="python bench.py --steps 80 --warmup 4 --no-cpu-baseline --e2e-steps 3"
timeout 300 $LONG > $out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_step -s 60 -c 1 -o $out/prof_step $LONG > $out/ncu_full.log 2>&1
echo "ncu full k_step rc=$?"
timeout 300 $SHORT > $out/plain3.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_observe -s 20 -c 1 -o $out/prof_observe $SHORT > $out/ncu_full_obs.log 2>&1
echo "ncu full k_observe rc=$?"
timeout 300 $SHORT > $out/plain4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_mt_step -s 60 -c 1 -o $out/prof_mt $SHORT > $out/ncu_full_mt.log 2>&1
echo "ncu full k_mt_step rc=$?"
timeout 300 $SHORT > $out/plain5.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_plan -s 3 -c 1 -o $out/prof_plan $SHORT > $out/ncu_full_plan.log 2>&1
echo "ncu full k_plan rc=$?"
timeout 300 python tools/plan_phase_clocks.py > $out/plan_phase.log 2>&1; cat $out/plan_phase.log
timeout 300 python tools/step_profile.py > $out/step_profile.txt 2>&1
timeout 300 python tools/mt_phase_clocks.py > $out/mt_phase.log 2>&1
ls -la $out
