#!/bin/bash
# Round-end measurement pass (run on the GPU box: `gpurun --timeout 900 -- bash tools/final_measure.sh`).
# Every ncu pass runs only after the same command has exited 0 without ncu; numbers printed under ncu are never bench values.
set -x
O=gpurun_out/f
mkdir -p $O
MINE='regex:conv_gemm_kernel|prep_input_kernel|nms_|rescale_merge_kernel|s2_gather_kernel|upsample2_kernel|add_kernel|copy_channels_kernel|maxpool_kernel'
timeout 300 python bench.py --steps 20 --warmup 3 > $O/bench.log 2> $O/bench.err; echo bench=$?
timeout 200 python tools/layer_times.py --batch 256 --reps 5 --counters --out $O/layer_times.txt > $O/lt.log 2>&1; echo lt=$?
# ablation floors per layer: full | no epilogue work (dbg 64) | no output stores (dbg 4)
timeout 200 python tools/layer_times.py --batch 256 --reps 5 --dbg 0,64,4 --out $O/layer_ablation.txt > $O/abl.log 2>&1; echo abl=$?
timeout 120 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/bench_short.log 2>&1; echo short=$?
# launch list of our kernels over the same command (cold-cache, serialised: shares, not absolutes)
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k "$MINE" -c 900 --csv \
  --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_launches.log 2>&1; echo ncu_list=$?
# DRAM traffic of every conv launch of one forward (the second forward of the run)
timeout 400 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:conv_gemm_kernel --launch-skip 75 -c 75 --csv --log-file $O/conv_traffic.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_traffic.log 2>&1; echo ncu_traffic=$?
# full capture of three conv launches: L3 (208^2 3x3 with residual), L14 (52^2 3x3 with residual), L64 (13^2 3x3)
for spec in "L3 3" "L14 11" "L64 45"; do
  set -- $spec
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_kernel --launch-skip $((75 + $2)) -c 1 \
    -f -o $O/conv_$1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_full_$1.log 2>&1; echo ncu_full_$1=$?
  ncu -i $O/conv_$1.ncu-rep --page raw --csv > $O/conv_$1.raw.csv 2>/dev/null
  python tools/ncu_extract.py $O/conv_$1.raw.csv > $O/conv_$1_ncu_full.txt
  ncu -i $O/conv_$1.ncu-rep --page source --csv > $O/conv_$1.source.csv 2>/dev/null
  ls -la $O/conv_$1.ncu-rep; rm -f $O/conv_$1.ncu-rep
done
tail -1 $O/bench.log
