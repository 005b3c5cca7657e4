#!/bin/bash
# round 2, call I: tile family with V from the half-transform's accumulators — parity, then the dense shapes
mkdir -p gpurun_out
python -m pytest tests -m "gpu and not slow" -q -x > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/i_pytest.log
b() { name=$1; shift; python bench.py --steps 8 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/i_$name.json 2> gpurun_out/i_$name.err; echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/i_$name.json") if l.startswith("{")][-1]); print(round(d["ms_per_step"],4), d["stage_ms"], d.get("gpu_launches_per_build_per_gpu"), (d.get("e2e") or {}).get("ms_per_step"))
except Exception as e: print("no line", e)
PY
)"; tail -2 gpurun_out/i_$name.err; }
b benzene --workload benzene_dimer
b benzene_old --workload benzene_dimer --opt coulomb_v_in_ht=0
b benzene_overlap --workload benzene_dimer --opt coulomb_overlap=1
b benzene_small_overlap --workload benzene_dimer --opt pp_bigreg=0 --opt exchange_producer=0 --opt coulomb_overlap=1
b w50dense --workload w50_dense --no-e2e
b water --workload water
b w50q600 --workload w50_q600 --no-e2e
python tools/sanitizer_cases.py > gpurun_out/i_cases.log 2>&1; tail -10 gpurun_out/i_cases.log
