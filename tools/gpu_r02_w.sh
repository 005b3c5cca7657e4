#!/bin/bash
# round 2, call W: what the driver runs at round end, on the final tree — the whole GPU suite (-x, slow test included), smoke(),
# the default bench line, the reference arm — then the ncu launch list of the bench command
mkdir -p gpurun_out
( time python -m pytest tests -x -q -m gpu ) > gpurun_out/w_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/w_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/w_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/w_smoke.log
( time python bench.py ) > gpurun_out/w_bench.json 2> gpurun_out/w_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/w_bench.err; head -c 300 gpurun_out/w_bench.json
( time python bench.py --impl reference --steps 20 --warmup 3 ) > gpurun_out/w_reference.json 2> gpurun_out/w_reference.err; echo "reference rc=$?"; tail -3 gpurun_out/w_reference.err; head -c 400 gpurun_out/w_reference.json
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/w_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/w_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/w_ncu_list.log 2>&1
echo "ncu list rc=$?"
