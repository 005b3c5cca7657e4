#!/bin/bash
# round 2, call O: what the driver runs at round end, on the final tree — the whole GPU suite (slow test included), smoke(), the
# default bench line and the reference arm
mkdir -p gpurun_out
( time python -m pytest tests -x -q -m gpu ) > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/o_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/o_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/o_smoke.log
( time python bench.py ) > gpurun_out/o_bench.json 2> gpurun_out/o_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/o_bench.err; head -c 300 gpurun_out/o_bench.json
( time python bench.py --impl reference --steps 20 --warmup 3 ) > gpurun_out/o_reference.json 2> gpurun_out/o_reference.err; echo "reference rc=$?"; tail -3 gpurun_out/o_reference.err; head -c 400 gpurun_out/o_reference.json
