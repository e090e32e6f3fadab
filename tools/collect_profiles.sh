#!/bin/bash
# Copy the judged summaries of the last tools/r2_measure.sh pass from gpurun_out/r2 (scratch) into profiles/ (tracked).
S=gpurun_out/r2; D=profiles
for c in c3 c1 c2 c5 f32x3 f32ffma reference; do [ -s $S/bench_$c.json ] && cp $S/bench_$c.json $D/r2_bench_$c.json; done
cp $S/layer_times.txt $D/r2_layer_times_b256.txt
cp $S/layer_ablation.txt $D/r2_layer_ablation_b256.txt
cp $S/attribution.txt $D/r2_attribution_pdl.txt
cp $S/attribution_nopdl.txt $D/r2_attribution_nopdl.txt
cp $S/launches.csv $D/r2_launches_b256.csv
cp $S/conv_traffic.csv $D/r2_conv_traffic.csv
cp $S/conv_traffic.json $D/r2_conv_traffic.json
for l in L3 L14 L64; do
  cp $S/conv_${l}_ncu_full.txt $D/r2_conv3x3_${l}_ncu_full.txt
  cp $S/conv_${l}_source_hotspots.txt $D/r2_conv3x3_${l}_source_hotspots.txt
done
ls -la $D | head -60
