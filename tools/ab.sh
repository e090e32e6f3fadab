#!/bin/bash
# A/B of two prebuilt libraries on the same box: bash tools/ab.sh A B A B
O=gpurun_out/ab; mkdir -p $O
for v in "$@"; do
  cp tools/ab/lib_$v.so amyloid_yolo_paper_b200/libamyloid_yolo_b200.so
  timeout 150 python tools/layer_times.py --batch 256 --reps 5 --out $O/lt_$v.txt > $O/lt_$v.log 2>&1
  echo "$v $(grep 'sum us' $O/lt_$v.txt) | $(grep 'whole forward' $O/lt_$v.txt)"
done
