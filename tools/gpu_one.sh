#!/bin/bash
# run a few test files quickly: tools/gpu_one.sh <tag> <pytest args...>
tag=$1; shift
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest "$@" -m gpu -x -q -s > $out/pytest.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|Error|error|skipped|assert|^E " $out/pytest.log | tail -15
