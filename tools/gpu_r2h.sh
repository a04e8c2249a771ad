#!/bin/bash
tag=${1:-r2_h}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_policy.py tests/test_reference_integration.py tests/test_mt.py -m gpu -x -q -s > $out/pytest_new.log 2>&1; echo "pytest rc=$?"
grep -E "policy rollout|mean reward|passed|failed|Error|error|skipped|assert" $out/pytest_new.log | tail -12
timeout 600 python tools/plan_profile.py > $out/plan_profile.txt 2>&1; cat $out/plan_profile.txt | cut -c1-330
