#!/bin/bash
# round 2, call H: checked build with the self-test that must fire; small shapes with the 9-warp kernels + Coulomb stream
mkdir -p gpurun_out
JCDF_LIBRARY=$PWD/juliachem.jl_b200/libjcdf_checked.so python tools/sanitizer_cases.py --fire > gpurun_out/h_checked_cases.log 2>&1; echo "cases rc=$?"; tail -4 gpurun_out/h_checked_cases.log
b() { name=$1; shift; python bench.py --steps 8 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/h_$name.json 2> gpurun_out/h_$name.err; echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/h_$name.json") if l.startswith("{")][-1]); print(round(d["ms_per_step"],4), d["stage_ms"], d.get("graph_replays"), (d.get("e2e") or {}).get("ms_per_step"))
except Exception as e: print("no line", e)
PY
)"; tail -2 gpurun_out/h_$name.err; }
S="--opt pp_bigreg=0 --opt exchange_producer=0"
b benzene --workload benzene_dimer
b benzene_small_kernels --workload benzene_dimer $S
b benzene_small_overlap --workload benzene_dimer $S --opt coulomb_fused=0 --opt coulomb_overlap=1
b benzene_small_overlap_nograph --workload benzene_dimer $S --opt coulomb_fused=0 --opt coulomb_overlap=1 --opt graph=0
b c20h42_small_overlap --workload c20h42 $S --opt coulomb_overlap=1
b w50_small_overlap $S --opt coulomb_overlap=1
