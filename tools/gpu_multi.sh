#!/bin/bash
# multi-GPU evidence: tools/gpu_multi.sh <tag> <N>
tag=$1; N=$2
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv > $out/gpus.txt
if [ "$N" -ge 2 ]; then
timeout 600 python -m pytest tests/test_nccl_two_ranks.py -m gpu -x -q > $out/pytest_nccl.log 2>&1; echo "nccl test rc=$?"; tail -3 $out/pytest_nccl.log
fi
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 tools/d2h_ceiling.py > $out/d2h_ceiling_$N.json 2> $out/d2h.err; echo "d2h rc=$?"; tail -1 $out/d2h_ceiling_$N.json
timeout 900 $TR --master-port 29512 bench.py --gpus $N --steps 200 --warmup 20 > $out/bench_$N.json 2> $out/bench_$N.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('$out/bench_$N.json').read().strip().splitlines()[-1])
    print('N=$N value %.4g e2e %.4g collectives %s' % (d['value'], d['e2e']['value'], json.dumps(d.get('collectives'))))
except Exception as e:
    print('no bench line', e)
PY
