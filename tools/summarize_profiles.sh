#!/bin/bash
# copy the evidence of one tools/gpu_check.sh call (gpurun_out/<tag>) into profiles/<tag>_* (run in the dev container)
tag=$1
src=gpurun_out/$tag
dst=profiles
cp $src/bench.json $dst/${tag}_bench.json
cp $src/bench_ref.json $dst/${tag}_bench_reference.json
cp $src/launches.csv $dst/${tag}_launches.csv
cp $src/pytest_gpu.log $dst/${tag}_pytest_gpu.txt
cp $src/phase.log $dst/${tag}_phase_clocks.txt
cp $src/plan_phase.log $dst/${tag}_plan_phase_clocks.txt
cp $src/mt_phase.log $dst/${tag}_mt_phase_clocks.txt
cp $src/step_profile.txt $dst/${tag}_step_profile.txt
for pair in prof_step:k_step prof_observe:k_observe prof_mt:k_mt prof_plan:k_plan_centroid; do
  rep=${pair%%:*}; name=${pair#*:}
  [ -f $src/$rep.ncu-rep ] || continue
  ncu -i $src/$rep.ncu-rep --page details > $dst/${tag}_${name}_full_details.txt 2>/dev/null
  ncu -i $src/$rep.ncu-rep --page raw --csv > $dst/${tag}_${name}_full_raw.csv 2>/dev/null
  ncu -i $src/$rep.ncu-rep --page source --csv --print-source cuda,sass > /tmp/${name}_source.csv 2>/dev/null
  python tools/ncu_lines.py /tmp/${name}_source.csv 120 > $dst/${tag}_${name}_source_lines.txt 2>/dev/null
done
ls -la $dst | grep $tag
