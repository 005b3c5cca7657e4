#!/bin/bash
# Scaling runs of bench.py on one 8 x B200 box, one process per GPU (the launch line of the bench contract).
#   gpurun --gpus 8 -- bash tools/scale_run.sh
mkdir -p gpurun_out
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/scale_w50_n$n.json 2> gpurun_out/scale_w50_n$n.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/scale_w50_n$n.json") if l.startswith("{")][-1])
    print("w50", d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["stage_ms"])
except Exception as e:
    print("w50 n=$n failed", e)
PY
  tail -2 gpurun_out/scale_w50_n$n.err
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
    bench.py --gpus 8 --workload gly18 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/scale_gly18_n8.json 2> gpurun_out/scale_gly18_n8.err
python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/scale_gly18_n8.json") if l.startswith("{")][-1])
    print("gly18", d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["stage_ms"], d["config"])
except Exception as e:
    print("gly18 failed", e)
PY
tail -3 gpurun_out/scale_gly18_n8.err
