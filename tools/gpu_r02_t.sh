#!/bin/bash
# round 2, call T: block-screened exchange on the chain-like benchmark shapes (where whole exchange tiles drop out)
mkdir -p gpurun_out
for a in "c20h42 0 10 20" "gly18_q1167 0 10 27"; do
  set -- $a
  ( time python tools/block_screen_cases.py $a ) > gpurun_out/t_block_screen_$1.jsonl 2> gpurun_out/t_block_screen_$1.err; echo "$1 rc=$?"
  python - <<PY
import json
for ln in open("gpurun_out/t_block_screen_$1.jsonl"):
    d = json.loads(ln)
    print(d["df_exchange_n_blocks"], d["blocks_kept"], d["blocks_total"], d["exchange_tiles_skipped"], d["exchange_tiles_total"], d["stage_ms_median_of_5"]["K"], d["stage_ms_median_of_5"]["total"], d["max_abs_err_kept_vs_default_build"], d["dropped_elements_equal_F_in_plus_H_bitwise"])
PY
  tail -3 gpurun_out/t_block_screen_$1.err
done
