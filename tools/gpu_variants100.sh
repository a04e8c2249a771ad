#!/bin/bash
# bench the 100v100 configuration with alternative builds (tools/variants/*.so)
tag=${1:-var100}
out=gpurun_out/$tag
mkdir -p $out
cp hmp2g_b200/libdca_b200.so /tmp/default.so
for v in tools/variants/*.so; do
  n=$(basename $v .so)
  cp $v hmp2g_b200/libdca_b200.so
  timeout 300 python bench.py --scenario 100v100 --no-cpu-baseline --no-other-configs --steps 360 --warmup 40 --e2e-steps 3 > $out/bench_$n.json 2> $out/bench_$n.err
  python - <<PY
import json
try:
    d=json.loads(open('$out/bench_$n.json').read().strip().splitlines()[-1])
    print('$n: value %.4g  pipelined ms/step %.4f  kernel_ms %.4f  frac %.3f  observe_ms %.4f' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['roofline_gather_kernel']['kernel_ms']))
except Exception as e:
    print('$n: no bench line', e)
PY
done
cp /tmp/default.so hmp2g_b200/libdca_b200.so
