#!/bin/bash
# quick GPU iteration: parity tests, bench line (short CPU baseline skipped), phase clocks
tag=${1:-quick}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 $out/pytest_gpu.log
timeout 600 python bench.py --no-cpu-baseline > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"
tail -3 $out/bench.err
python - <<PY
import json
try:
    d=json.loads(open('$out/bench.json').read().strip().splitlines()[-1])
    print('value %.4g  ms/step %.4f  kernel_ms %.4f  frac %.3f  e2e %.4g' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['e2e']['value']))
except Exception as e:
    print('no bench line', e)
PY
timeout 300 python tools/phase_clocks.py > $out/phase.log 2>&1; cat $out/phase.log
