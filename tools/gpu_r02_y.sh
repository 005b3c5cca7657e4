#!/bin/bash
# round 2, call Y: after the 64-row half-tile skipping of the block-screened exchange — the chain shapes again, then the round-end
# sequence on the final tree (whole GPU suite -x, smoke, default bench line)
mkdir -p gpurun_out
for a in "c20h42 0 10" "gly18_q1167 0 10"; do
  set -- $a
  python tools/block_screen_cases.py $a > gpurun_out/y_block_screen_$1.jsonl 2> gpurun_out/y_block_screen_$1.err; echo "$1 rc=$?"
  python - <<PY
import json
for ln in open("gpurun_out/y_block_screen_$1.jsonl"):
    d = json.loads(ln)
    print(d["df_exchange_n_blocks"], d["blocks_kept"], d["blocks_total"], d["exchange_tiles_skipped"], d["exchange_half_tiles_skipped"], d["exchange_tiles_total"], d["stage_ms_median_of_5"]["K"], d["max_abs_err_kept_vs_default_build"], d["dropped_elements_equal_F_in_plus_H_bitwise"])
PY
  tail -2 gpurun_out/y_block_screen_$1.err
done
( time python -m pytest tests -x -q -m gpu ) > gpurun_out/y_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/y_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/y_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/y_smoke.log
python bench.py > gpurun_out/y_bench.json 2> gpurun_out/y_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/y_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/y_bench.json"))
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["stage_ms"], d["parity_check"]["max_abs_err"], d["parity_check"].get("max_abs_err_vs_oracle"), d["roofline"]["frac"], d["cpu_baseline"]["value"])
PY
