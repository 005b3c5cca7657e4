#!/bin/bash
# round 2, call P: the no-allocation entry point — its test alone in a fresh process, then the whole GPU suite (the test runs last)
mkdir -p gpurun_out
( time python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu -s ) > gpurun_out/p_no_alloc_alone.log 2>&1; echo "alone rc=$?"; tail -12 gpurun_out/p_no_alloc_alone.log
python tools/no_alloc_timing.py > gpurun_out/p_no_alloc_timing.jsonl 2> gpurun_out/p_no_alloc_timing.err; echo "timing rc=$?"; cat gpurun_out/p_no_alloc_timing.jsonl; tail -3 gpurun_out/p_no_alloc_timing.err
( time python -m pytest tests -x -q -m gpu ) > gpurun_out/p_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/p_pytest.log
