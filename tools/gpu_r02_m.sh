#!/bin/bash
# round 2, call M (8 GPUs): the whole GPU test suite on a multi-GPU box (incl. the multi-process seam), then the scaling run of
# the final code: w50 at 1/2/4/8 ranks, gly18 at 8
mkdir -p gpurun_out
python -m pytest tests -m "gpu and not slow" -q -rs > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/m_pytest.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/m_w50_n1.json 2> gpurun_out/m_w50_n1.err; tail -c 200 gpurun_out/m_w50_n1.json
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/m_w50_n$n.json 2> gpurun_out/m_w50_n$n.err
done
for n in 1 2 4 8; do python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/m_w50_n$n.json") if l.startswith("{")][-1])
    print("w50", d["n_gpus"], round(d["value"],3), round(d["ms_per_step"],4), round(d["e2e"]["value"],3), d["stage_ms"], d["parity_check"]["max_abs_err"])
except Exception as e:
    print("w50 n=$n failed", e)
PY
done
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
    bench.py --gpus 8 --workload gly18 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/m_gly18_n8.json 2> gpurun_out/m_gly18_n8.err
python - <<PY
import json
try:
    d = json.loads([l for l in open("gpurun_out/m_gly18_n8.json") if l.startswith("{")][-1])
    print("gly18", d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["stage_ms"])
except Exception as e:
    print("gly18 failed", e)
PY
# which NCCL algorithm serves the 12.5 MB all-reduce at 8 ranks (NVLS = in-switch reduction)
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL,TUNING timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 \
    bench.py --gpus 8 --steps 3 --warmup 3 --no-e2e > gpurun_out/m_nccl_debug.json 2> gpurun_out/m_nccl_debug.err
grep -E "NVLS|Using network|Channel|algorithm|Algo|nvls" gpurun_out/m_nccl_debug.err | sort | uniq -c | sort -rn | head -12 > gpurun_out/m_nccl_summary.txt; cat gpurun_out/m_nccl_summary.txt | cut -c1-200
