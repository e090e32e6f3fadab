for D in 0 1 2 3 4 8 12 7; do
AY_DBG=$D timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"conv_gemm|first_conv|s2_gather|upsample2" -c 82 --csv --log-file gpurun_out/abl_$D.csv python bench.py --steps 1 --warmup 1 --batch 64 --no-cpu-baseline > gpurun_out/abl_$D.log 2>&1
done
