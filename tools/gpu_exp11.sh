out=gpurun_out/exp11; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "matches_reference_golden or no_resync or full_size_4096 or 100v100 or ragged or overlapping" > $out/pytest.log 2>&1; echo "parity subset rc=$? $(tail -1 $out/pytest.log)"
timeout 300 python bench.py --no-cpu-baseline --steps 900 --warmup 100 --e2e-steps 3 > $out/bench.json 2>$out/bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/exp11/bench.json').read().strip().splitlines()[-1])
print('value %.4g ms/step %.4f kernel_ms %.4f gather %.4f' % (d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['roofline_gather_kernel']['kernel_ms']))
PY
for st in 110; do echo "== step $st"; timeout 300 python tools/cta_timeline.py $st 2>&1 | grep -E "event time|lifetime|^mean|^env" | cut -c1-130; done
