#!/bin/bash
# round-2 call B: allocation-free entry point test (alone, then after the others), dram range captures
tag=${1:-r2_b}
out=gpurun_out/$tag
mkdir -p $out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 $out/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_gpu.log
for k in observe step; do
  R="python tools/dram_range.py --kernel $k --launches 64"
  timeout 300 $R > $out/range_${k}_plain.log 2>&1 &&
  timeout 600 ncu --replay-mode range --cache-control none --clock-control none \
     --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv --log-file $out/range_$k.csv $R > $out/range_${k}_ncu.log 2>&1
  echo "range $k rc=$?"; tail -1 $out/range_${k}_plain.log; tail -5 $out/range_$k.csv; tail -3 $out/range_${k}_ncu.log
  if [ ! -s $out/range_$k.csv ] || ! grep -q dram__bytes $out/range_$k.csv; then
    timeout 900 ncu --replay-mode app-range --cache-control none --clock-control none \
     --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv --log-file $out/range_$k.csv $R > $out/range_${k}_ncu2.log 2>&1
    echo "app-range $k rc=$?"; tail -5 $out/range_$k.csv; tail -3 $out/range_${k}_ncu2.log
  fi
done
ls -la $out
