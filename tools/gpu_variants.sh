#!/bin/bash
# bench alternative builds of libdca_b200.so (tools/variants/*.so) after the default one
tag=${1:-var}
out=gpurun_out/$tag
mkdir -p $out
cp hmp2g_b200/libdca_b200.so /tmp/default.so
for v in tools/variants/*.so; do
  n=$(basename $v .so)
  cp $v hmp2g_b200/libdca_b200.so
  timeout 300 python bench.py --no-cpu-baseline --no-other-configs --steps 900 --warmup 100 --e2e-steps 3 > $out/bench_$n.json 2> $out/bench_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('$out/bench_$n.json').read().strip().splitlines()[-1])
    print('$n: value %.4g  kernel_ms %.4f  frac %.3f  sync_start %.4f  observe_ms %.4f  scripted %.4f' % (d['value'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['roofline'].get('kernel_ms_synchronised_start', 0), d['roofline_gather_kernel']['kernel_ms'], d['with_scripted_planner']['ms_per_step']))
except Exception as e:
    print('$n: no bench line', e)
PY
done
cp /tmp/default.so hmp2g_b200/libdca_b200.so
