#!/bin/bash
# round 2, call X (8 GPUs, one process per GPU): the final tree at 8 ranks — w50 with the fixture parity on every rank, then the
# gly18 shape with and without the block-screened exchange (df_exchange_screen at the reference's default 10 K blocks)
mkdir -p gpurun_out
run() {  # name, extra args
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) \
      bench.py --gpus 8 "$@" > gpurun_out/x_$name.json 2> gpurun_out/x_$name.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/x_$name.json") if l.startswith("{")][-1])
    print("$name", d["n_gpus"], round(d["value"], 3), round(d["ms_per_step"], 3), d["e2e"] and round(d["e2e"]["value"], 3), d["stage_ms"], d["parity_check"] and d["parity_check"]["max_abs_err"], d["config"].get("exchange_block_screen"))
except Exception as e:
    print("$name failed", e)
PY
  tail -2 gpurun_out/x_$name.err
}
run w50_n8 --steps 10 --warmup 3
run gly18_n8 --workload gly18 --steps 5 --warmup 3 --no-cpu-baseline
run gly18_n8_xs10 --workload gly18 --steps 5 --warmup 3 --no-cpu-baseline --opt exchange_block_screen=10
