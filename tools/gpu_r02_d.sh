#!/bin/bash
# round 2, call D (8 GPUs): multi-device tests, w50 at 2/4/8 ranks (one process per GPU), gly18 at 8, single-process 8-device run
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/d_topo.txt 2>&1
python -m pytest tests -m "gpu and not slow" -q -rs -k "multi_device or ndev or s10 or water_scf or sharding or two_shards" > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/d_pytest.log
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/d_w50_n$n.json 2> gpurun_out/d_w50_n$n.err
  python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/d_w50_n$n.json") if l.startswith("{")][-1])
    print("w50", d["n_gpus"], round(d["value"],3), round(d["ms_per_step"],4), d["e2e"]["value"], d["e2e"]["gap_breakdown"], d["stage_ms"], d["parity_check"], d["rank0_cpu_affinity"])
except Exception as e:
    print("w50 n=$n failed", e)
PY
  tail -2 gpurun_out/d_w50_n$n.err
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
    bench.py --gpus 8 --workload gly18 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/d_gly18_n8.json 2> gpurun_out/d_gly18_n8.err
python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/d_gly18_n8.json") if l.startswith("{")][-1])
    print("gly18", d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["stage_ms"])
except Exception as e:
    print("gly18 failed", e)
PY
tail -3 gpurun_out/d_gly18_n8.err
python bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/d_w50_single_process_8dev.json 2> gpurun_out/d_w50_sp8.err; tail -c 700 gpurun_out/d_w50_single_process_8dev.json; tail -2 gpurun_out/d_w50_sp8.err
