out=gpurun_out/exp10; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $out/pytest_gpu.log
timeout 300 python bench.py --no-cpu-baseline --steps 900 --warmup 100 --e2e-steps 3 > $out/bench.json 2>$out/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/exp10/bench.json').read().strip().splitlines()[-1])
print('value %.4g ms/step %.4f kernel_ms %.4f gather %.4f' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline_gather_kernel']['kernel_ms']))
PY
for st in 25 110; do echo "== step $st"; timeout 300 python tools/cta_timeline.py $st 2>&1 | grep -E "event time|lifetime|replacement|^mean|^env|done envs"; done
timeout 300 python tools/step_profile.py 2>&1 | head -20
