#!/bin/bash
# round 2, last call: the new full-size block-screened test and one short run of the final bench.py (its traffic constants changed)
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_fullsize.py -x -q -m gpu -s -k block_screened ) > gpurun_out/zz_fullsize_block_screen.log 2>&1; echo "rc=$?"; grep "block-screened\|passed\|failed" gpurun_out/zz_fullsize_block_screen.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/zz_bench.json 2> gpurun_out/zz_bench.err; echo "bench rc=$?"; tail -2 gpurun_out/zz_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/zz_bench.json"))
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["traffic"], d["roofline"]["traffic_source"], d["parity_check"]["max_abs_err"])
PY
