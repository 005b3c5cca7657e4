#!/bin/bash
# round 2, call B: option A/B timings (work-list order, graph, exchange slices at the 8-GPU shard size, Coulomb overlap on the
# small shapes), exchange CTA balance at Q_g = 600, and compute-sanitizer memcheck over the smallest case of every family
mkdir -p gpurun_out
b() { name=$1; shift; python bench.py --steps 6 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err; echo "$name rc=$? $(python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/b_$name.json") if l.startswith("{")][-1]); print(round(d["ms_per_step"],4), d["stage_ms"], d.get("graph_replays"), (d.get("e2e") or {}).get("ms_per_step"))
except Exception as e: print("no line", e)
PY
)"; }
b w50_default
b w50_pmajor --opt pp_order=1
b w50_nograph --opt graph=0
b w50_9warps --opt pp_bigreg=0
b w50_xinwarp --opt exchange_producer=0
b c20h42_9warps --workload c20h42 --opt pp_bigreg=0
for s in 24 33 48 64; do b q600_s$s --workload w50_q600 --no-e2e --opt exchange_slices=$s; done
b q600_auto --workload w50_q600 --no-e2e
b benzene --workload benzene_dimer
b benzene_overlap --workload benzene_dimer --opt coulomb_overlap=1
b benzene_nograph --workload benzene_dimer --opt graph=0
b benzene_9warps --workload benzene_dimer --opt pp_bigreg=0
b c20h42 --workload c20h42
b c20h42_overlap --workload c20h42 --opt coulomb_overlap=1
b water --workload water
b water_nograph --workload water --opt graph=0
python tools/exchange_balance.py w50_q600 > gpurun_out/b_balance_q600.log 2>&1; tail -5 gpurun_out/b_balance_q600.log
python tools/sanitizer_cases.py > gpurun_out/b_cases_plain.log 2>&1 && \
compute-sanitizer --tool memcheck --log-file gpurun_out/b_memcheck.log python tools/sanitizer_cases.py > gpurun_out/b_memcheck_run.log 2>&1
echo "memcheck rc=$?"; tail -4 gpurun_out/b_memcheck.log; tail -9 gpurun_out/b_memcheck_run.log
