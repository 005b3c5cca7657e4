#!/bin/bash
# round 2, call C: item records + p-major default; small shapes with the two-pass Coulomb on a second stream; tests
mkdir -p gpurun_out
b() { name=$1; shift; python bench.py --steps 6 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/c_$name.json 2> gpurun_out/c_$name.err; echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/c_$name.json") if l.startswith("{")][-1]); print(round(d["ms_per_step"],4), d["stage_ms"], d.get("graph_replays"), (d.get("e2e") or {}).get("ms_per_step"))
except Exception as e: print("no line", e)
PY
)"; tail -2 gpurun_out/c_$name.err; }
b w50
b w50_9warps --opt pp_bigreg=0
b c20h42 --workload c20h42
b c20h42_2x4 --workload c20h42 --opt pp_layout=0
b gly18q --workload gly18_q1167 --no-e2e
b benzene --workload benzene_dimer
b benzene_2pass --workload benzene_dimer --opt coulomb_fused=0
b benzene_2pass_overlap --workload benzene_dimer --opt coulomb_fused=0 --opt coulomb_overlap=1
b benzene_fused --workload benzene_dimer --opt coulomb_fused=1
b w50dense --workload w50_dense --no-e2e
b w50dense_9warps --workload w50_dense --no-e2e --opt pp_bigreg=0
python -m pytest tests -m "gpu and not slow" -q -x > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/c_pytest.log
