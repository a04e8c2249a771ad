#!/bin/bash
# planner iteration: planner parity tests (+ the scripted-planner scale test), phase clocks, ncu full of k_plan
tag=${1:-plan}
out=gpurun_out/$tag
mkdir -p $out
if [ "$3" != "notest" ]; then
timeout 900 python -m pytest tests/test_planner.py tests/test_gpu_parity_scale.py -m gpu -x -q -s > $out/pytest_plan.log 2>&1; echo "pytest rc=$?"
grep -E "scripted|passed|failed|Error|error" $out/pytest_plan.log | tail -8
fi
timeout 300 python tools/plan_phase_clocks.py > $out/plan_phase.log 2>&1; cat $out/plan_phase.log
timeout 300 python bench.py --planner scripted --steps 360 --warmup 20 --no-cpu-baseline --no-other-configs --e2e-steps 3 > $out/bench_scripted.json 2> $out/bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('$out/bench_scripted.json').read().strip().splitlines()[-1])
    print('scripted: value %.4g  ms/step %.4f planner_kernel_ms %s' % (d['value'], d['ms_per_step'], d['config'].get('planner_kernel_ms')))
except Exception as e:
    print('no bench line', e)
PY
if [ "$2" = "ncu" ]; then
SHORT="python bench.py --steps 12 --warmup 4 --no-cpu-baseline --no-other-configs --e2e-steps 3"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_plan$ -s 30 -c 1 -o $out/prof_plan $SHORT --planner scripted > $out/ncu_full_plan.log 2>&1
echo "ncu full k_plan rc=$?"
fi
