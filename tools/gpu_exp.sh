#!/bin/bash
# experiment call: GPU parity tests on the default build, bench with / without planner overlap, then the variants
tag=${1:-exp}
out=gpurun_out/$tag
mkdir -p $out
if [ -z "$SKIP_PYTEST" ]; then
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 $out/pytest_gpu.log
fi
B="--no-cpu-baseline --steps 900 --warmup 100 --e2e-steps 3"
show() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print('%s: value %.4g  ms/step %.4f  kernel_ms %.4f  frac %.3f  gather_ms %.4f' % (sys.argv[1], d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['roofline_gather_kernel']['kernel_ms']))
except Exception as e:
    print(sys.argv[1], 'no bench line', e)
PY
}
timeout 300 python bench.py $B > $out/bench_default.json 2> $out/bench_default.err; show default $out/bench_default.json
cp hmp2g_b200/libdca_b200.so /tmp/default.so
shopt -s nullglob
for v in tools/variants/*.so; do
  n=$(basename $v .so)
  cp $v hmp2g_b200/libdca_b200.so
  [[ $n == diag* ]] || timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "matches_reference_golden or no_resync or full_size_4096 or 100v100 or ragged" > $out/pytest_$n.log 2>&1; echo "$n parity subset rc=$? $(tail -1 $out/pytest_$n.log)"
  timeout 300 python bench.py $B > $out/bench_$n.json 2> $out/bench_$n.err; show $n $out/bench_$n.json
done
cp /tmp/default.so hmp2g_b200/libdca_b200.so
timeout 300 python tools/phase_clocks.py > $out/phase.log 2>&1; cat $out/phase.log
