#!/bin/bash
# round 2, call R: per-function half-transform with two CTAs per SM (pp_layout = 2) — parity, then A/B bench lines on one box
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "warp_layouts or per_function or register_split or refused" ) > gpurun_out/r_layout_tests.log 2>&1; echo "layout tests rc=$?"; tail -8 gpurun_out/r_layout_tests.log
for o in "pp_layout=0" "pp_layout=2"; do
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --opt $o > gpurun_out/r_bench_w50_${o/=/}.json 2> gpurun_out/r_bench_w50_${o/=/}.err; echo "w50 $o rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r_bench_w50_${o/=/}.json"))
print("$o", d["ms_per_step"], d["stage_ms"], d["parity_check"]["max_abs_err"])
PY
done
for wl in c20h42 gly18_q1167; do for o in "pp_layout=-1" "pp_layout=2"; do
  python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --opt $o > gpurun_out/r_bench_${wl}_${o/=/}.json 2> gpurun_out/r_bench_${wl}_${o/=/}.err; echo "$wl $o rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r_bench_${wl}_${o/=/}.json"))
print("$wl $o", d["ms_per_step"], d["stage_ms"])
PY
done; done
