#!/bin/bash
# Round-2 measurement pass (run on the GPU box: `gpurun --timeout 1800 -- bash tools/r2_measure.sh`).
# Every ncu pass runs only after the same command has exited 0 without ncu; numbers printed under ncu are never bench values.
set -x
O=gpurun_out/r2
mkdir -p $O
MINE='regex:conv_gemm_kernel|prep_input_kernel|nms_|rescale_merge_kernel|s2_gather_kernel|upsample2_kernel|add_kernel|copy_channels_kernel|maxpool_kernel'
SHORT="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 400 python bench.py --steps 20 --warmup 3 > $O/bench_c3.json 2> $O/bench_c3.err; echo c3=$?
timeout 300 python bench.py --config c1 --steps 50 --warmup 5 > $O/bench_c1.json 2> $O/bench_c1.err; echo c1=$?
timeout 300 python bench.py --config c2 --steps 20 --warmup 3 > $O/bench_c2.json 2> $O/bench_c2.err; echo c2=$?
timeout 400 python bench.py --config c5 --steps 5 --warmup 3 > $O/bench_c5.json 2> $O/bench_c5.err; echo c5=$?
timeout 300 python bench.py --precision fp32 --batch 64 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $O/bench_f32x3.json 2> $O/f32x3.err; echo f32x3=$?
AY_F32_FFMA=1 timeout 600 python bench.py --precision fp32 --batch 64 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $O/bench_f32ffma.json 2> $O/f32ffma.err; echo ffma=$?
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo ref=$?
timeout 200 python tools/layer_times.py --batch 256 --reps 5 --counters --out $O/layer_times.txt > $O/lt.log 2>&1; echo lt=$?
timeout 200 python tools/layer_times.py --batch 256 --reps 5 --dbg 0,64,4 --out $O/layer_ablation.txt > $O/abl.log 2>&1; echo abl=$?
timeout 200 python tools/attribution.py --out $O/attribution.txt > $O/attr.log 2>&1; echo attr=$?
AY_NO_PDL=1 timeout 200 python tools/attribution.py --out $O/attribution_nopdl.txt > $O/attr_nopdl.log 2>&1; echo attr_nopdl=$?
timeout 120 $SHORT > $O/bench_short.log 2>&1; echo short=$?
# launch list of our kernels over the same command (cold-cache, serialised: shares, not absolutes)
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -k "$MINE" -c 1200 --csv \
  --log-file $O/launches.csv $SHORT > $O/ncu_launches.log 2>&1; echo ncu_list=$?
# DRAM traffic of every conv launch of one forward (the second forward of the run)
timeout 500 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:conv_gemm_kernel --launch-skip 75 -c 75 --csv --log-file $O/conv_traffic.csv \
  $SHORT > $O/ncu_traffic.log 2>&1; echo ncu_traffic=$?
python tools/traffic_json.py $O/conv_traffic.csv 256 > $O/conv_traffic.json
# full capture of three conv launches: L3 (208^2 3x3 with residual), L14 (52^2 3x3 with residual), L64 (13^2 3x3)
for spec in "L3 3" "L14 11" "L64 45"; do
  set -- $spec
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_kernel --launch-skip $((75 + $2)) -c 1 \
    -f -o $O/conv_$1 $SHORT > $O/ncu_full_$1.log 2>&1; echo ncu_full_$1=$?
  ncu -i $O/conv_$1.ncu-rep --page raw --csv > $O/conv_$1.raw.csv 2>/dev/null
  python tools/ncu_extract.py $O/conv_$1.raw.csv > $O/conv_$1_ncu_full.txt
  ncu -i $O/conv_$1.ncu-rep --page source --csv > $O/conv_$1.source.csv 2>/dev/null
  python tools/src_hot.py $O/conv_$1.source.csv > $O/conv_$1_source_hotspots.txt 2>/dev/null
  ls -la $O/conv_$1.ncu-rep; rm -f $O/conv_$1.ncu-rep $O/conv_$1.source.csv
done
tail -c 600 $O/bench_c3.json
