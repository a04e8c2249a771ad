#!/bin/bash
# full check: GPU test suite, bench (default), driver-like short bench of both arms
tag=${1:-full}
out=gpurun_out/$tag
mkdir -p $out
timeout 1800 python -m pytest tests -m gpu -x -q -s > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
grep -E "policy rollout|mean reward|passed|failed|Error|error|skipped" $out/pytest_gpu.log | tail -12
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; tail -3 $out/bench.err
timeout 600 python bench.py --steps 20 --warmup 5 > $out/bench_driverlike.json 2>> $out/bench.err; echo "bench driver-like rc=$?"
timeout 400 python bench.py --impl reference --steps 20 --warmup 5 > $out/bench_ref.json 2>> $out/bench.err; echo "ref rc=$?"
python - <<PY
import json
for f in ('bench.json','bench_driverlike.json'):
    try:
        d=json.loads(open('$out/'+f).read().strip().splitlines()[-1])
        r=d['roofline']
        print(f,'value %.4g ms/step %.4f kernel_ms %.4f frac %.3f ep_avg %.4f sync_start %.4f dram_frac %s gather frac %.3f dram %s e2e %.4g' % (d['value'], d['ms_per_step'], r['kernel_ms'], r['frac'], r['kernel_ms_episode_avg'], r['kernel_ms_synchronised_start'], r['dram_frac'], d['roofline_gather_kernel']['frac'], d['roofline_gather_kernel']['dram_frac'], d['e2e']['value']))
        print('  scripted', d['with_scripted_planner']['ms_per_step'], d['with_scripted_planner']['k_plan_ms'])
        print('  policy', json.dumps(d.get('policy_rollout_50v55')))
    except Exception as e:
        print(f,'no bench line', e)
PY
