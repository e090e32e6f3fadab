#!/bin/bash
# Round-2 first GPU pass: smoke, full GPU test suite, the four bench configs, attribution with and without PDL,
# per-layer ablations.  Every command under its own timeout; logs under gpurun_out/c1/.
O=gpurun_out/c1
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/gpu.txt 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo smoke=$?
if ! grep -q "smoke ok" $O/smoke.log; then
  echo "smoke failed with PDL; retrying without"; tail -5 $O/smoke.log
  export AY_NO_PDL=1
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke_nopdl.log 2>&1; echo smoke_nopdl=$?
  tail -3 $O/smoke_nopdl.log
fi
timeout 1500 python -m pytest tests -m gpu -q -s > $O/tests.log 2>&1; echo tests=$?
tail -15 $O/tests.log
timeout 400 python bench.py --steps 10 --warmup 3 > $O/bench_c3.json 2> $O/bench_c3.err; echo c3=$?
timeout 300 python bench.py --config c1 --steps 30 --warmup 5 > $O/bench_c1.json 2> $O/bench_c1.err; echo c1=$?
timeout 300 python bench.py --config c2 --steps 10 --warmup 3 > $O/bench_c2.json 2> $O/bench_c2.err; echo c2=$?
timeout 400 python bench.py --config c5 --steps 5 --warmup 3 > $O/bench_c5.json 2> $O/bench_c5.err; echo c5=$?
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo ref=$?
timeout 200 python tools/attribution.py --out $O/attribution_pdl.txt > $O/attr_pdl.log 2>&1; echo attr=$?
AY_NO_PDL=1 timeout 200 python tools/attribution.py --out $O/attribution_nopdl.txt > $O/attr_nopdl.log 2>&1; echo attr_nopdl=$?
timeout 300 python tools/layer_times.py --batch 256 --reps 4 --dbg 0,1,2,64 --counters --out $O/layer_times.txt > $O/lt.log 2>&1; echo lt=$?
AY_NO_HEADFAST=1 timeout 200 python tools/layer_times.py --batch 256 --reps 4 --out $O/layer_times_noheadfast.txt > $O/lt2.log 2>&1; echo lt_noheadfast=$?
for f in $O/bench_c*.json $O/bench_ref.json; do echo "== $f"; tail -c 1800 $f; echo; done
tail -12 $O/attribution_pdl.txt; tail -12 $O/attribution_nopdl.txt
grep -E "sum us|whole forward" $O/layer_times.txt $O/layer_times_noheadfast.txt
