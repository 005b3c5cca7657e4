#!/bin/bash
# round 2, call F: deferred W stores in the per-function half-transform — parity first, then timings
mkdir -p gpurun_out
python -m pytest tests -m "gpu and not slow" -q -x > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/f_pytest.log
b() { name=$1; shift; python bench.py --steps 6 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/f_$name.json 2> gpurun_out/f_$name.err; echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/f_$name.json") if l.startswith("{")][-1]); print(round(d["ms_per_step"],4), d["stage_ms"], d.get("graph_replays"), (d.get("e2e") or {}).get("ms_per_step"), d.get("parity_check"))
except Exception as e: print("no line", e)
PY
)"; tail -2 gpurun_out/f_$name.err; }
b w50
b c20h42 --workload c20h42
b gly18q --workload gly18_q1167 --no-e2e
b w50_q600 --workload w50_q600 --no-e2e
