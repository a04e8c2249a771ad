#!/bin/bash
out=gpurun_out/exp5; mkdir -p $out
B="--no-cpu-baseline --steps 900 --warmup 100 --e2e-steps 3"
for v in "" "--no-overlap"; do
timeout 300 python bench.py $B --planner scripted $v > $out/bench_scripted.json 2> $out/bench_scripted.err
python - "$v" <<'PY'
import json,sys
d=json.loads(open('gpurun_out/exp5/bench_scripted.json').read().strip().splitlines()[-1])
print('scripted %s: value %.4g ms/step %.4f kernel_ms %.4f planner_ms %.4f alive %.3f done %.3f' % (sys.argv[1], d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['config']['planner_kernel_ms'], d['config']['alive_fraction_at_end'], d['config']['done_fraction_last_step']))
PY
done
timeout 300 python tools/plan_phase_clocks.py 2>&1 | tail -12
