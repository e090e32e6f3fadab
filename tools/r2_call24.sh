#!/bin/bash
bash tools/ab.sh A B A B A B
for v in A B; do echo "== $v"; grep -E "^ *(3|7|10|14|17|20|35) " gpurun_out/ab/lt_$v.txt | cut -c1-44; done
cp tools/ab/lib_B.so amyloid_yolo_paper_b200/libamyloid_yolo_b200.so
