#!/bin/bash
# round 2, call U (2 GPUs): the whole GPU suite where the 2-device cases run too (block-screened exchange on 2 devices among
# them), then block-screened bench lines, one process per GPU
mkdir -p gpurun_out
( time python -m pytest tests -x -q -m gpu ) > gpurun_out/u_pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/u_pytest_2gpu.log
for o in 0 10; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload c20h42 --steps 10 --warmup 3 --opt exchange_block_screen=$o > gpurun_out/u_bench_c20h42_n2_xs$o.json 2> gpurun_out/u_bench_c20h42_n2_xs$o.err; echo "c20h42 n2 xs=$o rc=$?"
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/u_bench_c20h42_n2_xs$o.json") if l.startswith("{")][-1])
print(d["value"], d["ms_per_step"], d["stage_ms"], d["config"].get("exchange_block_screen"), d["e2e"]["value"])
PY
done
python bench.py --workload c20h42 --steps 10 --warmup 3 --no-cpu-baseline --opt exchange_block_screen=10 > gpurun_out/u_bench_c20h42_n1_xs10.json 2> gpurun_out/u_bench_c20h42_n1_xs10.err; echo "c20h42 n1 xs=10 rc=$?"; head -c 250 gpurun_out/u_bench_c20h42_n1_xs10.json
python bench.py --steps 5 --warmup 3 --no-cpu-baseline --opt exchange_block_screen=50 > gpurun_out/u_bench_w50_n1_xs50.json 2> gpurun_out/u_bench_w50_n1_xs50.err; echo "w50 n1 xs=50 rc=$?"; head -c 250 gpurun_out/u_bench_w50_n1_xs50.json; tail -2 gpurun_out/u_bench_w50_n1_xs50.err
