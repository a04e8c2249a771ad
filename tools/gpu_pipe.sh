#!/bin/bash
tag=${1:-pipe}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_parity_scale.py -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $out/pytest.log
timeout 300 python tools/pipe_bench.py > $out/pipe_bench.txt 2>&1; cat $out/pipe_bench.txt
