#!/bin/bash
# round 2: bench lines of the final tree for the C20H42 shape with the block-screened exchange (tiles + half tiles) and the real benzene dimer
mkdir -p gpurun_out
python bench.py --workload c20h42 --steps 10 --warmup 3 --no-cpu-baseline --opt exchange_block_screen=10 > gpurun_out/zzz_c20h42_xs10.json 2> gpurun_out/zzz_c20h42_xs10.err; echo "rc=$?"
python bench.py --workload benzene_dimer_real --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/zzz_benzene_real.json 2> gpurun_out/zzz_benzene_real.err; echo "rc=$?"
python - <<'PY'
import json
for f in ("c20h42_xs10", "benzene_real"):
    d = json.load(open("gpurun_out/zzz_%s.json" % f))
    print(f, round(d["value"], 2), round(d["ms_per_step"], 3), d["e2e"] and round(d["e2e"]["value"], 2), d["stage_ms"], d["config"].get("exchange_block_screen"), d["parity_check"])
PY
