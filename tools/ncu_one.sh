#!/bin/bash
# one `ncu --set full` capture of the k-th conv launch of the second forward: bash tools/ncu_one.sh NAME K
O=gpurun_out/f; mkdir -p $O
timeout 120 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/pre_$1.log 2>&1 || exit 1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_kernel --launch-skip $((75 + $2)) -c 1 \
  -f -o $O/conv_$1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_full_$1.log 2>&1; echo ncu_full_$1=$?
ncu -i $O/conv_$1.ncu-rep --page raw --csv > $O/conv_$1.raw.csv 2>/dev/null
python tools/ncu_extract.py $O/conv_$1.raw.csv > $O/conv_$1_ncu_full.txt
ncu -i $O/conv_$1.ncu-rep --page source --csv > $O/conv_$1.source.csv 2>/dev/null
rm -f $O/conv_$1.ncu-rep
