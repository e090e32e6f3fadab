#!/bin/bash
# ncu --set full of the NMS mask kernel at the stress size (config C5), after the same command has exited 0 without ncu
O=gpurun_out/c30; mkdir -p $O
CMD="python bench.py --config c5 --steps 1 --warmup 3 --no-e2e --no-cpu-baseline"
timeout 300 $CMD > $O/pre.log 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:nms_mask_kernel --launch-skip 3 -c 1 -f -o $O/nms_mask $CMD > $O/ncu.log 2>&1; echo ncu=$?
ncu -i $O/nms_mask.ncu-rep --page raw --csv > $O/nms_mask.raw.csv 2>/dev/null
python tools/ncu_extract.py $O/nms_mask.raw.csv > $O/nms_mask_ncu_full.txt
rm -f $O/nms_mask.ncu-rep
head -40 $O/nms_mask_ncu_full.txt
