#!/bin/bash
# round 2, call Q: block-screened exchange (df_exchange_screen) — its tests, the parity suite of the kernels it touches, the real w50 record
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_block_screened_exchange.py -x -q -m gpu ) > gpurun_out/q_block_screen_tests.log 2>&1; echo "block-screen tests rc=$?"; tail -15 gpurun_out/q_block_screen_tests.log
( time python -m pytest tests/test_gpu_parity.py tests/test_gpu_water_scf.py -x -q -m gpu ) > gpurun_out/q_parity.log 2>&1; echo "parity rc=$?"; tail -6 gpurun_out/q_parity.log
( time python tools/block_screen_cases.py 0 25 50 ) > gpurun_out/q_block_screen_w50.jsonl 2> gpurun_out/q_block_screen_w50.err; echo "w50 rc=$?"; cat gpurun_out/q_block_screen_w50.jsonl; tail -5 gpurun_out/q_block_screen_w50.err
