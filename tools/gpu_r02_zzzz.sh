#!/bin/bash
# round 2: the checked build (bounds checks) over the block-screened cases that skip tiles and half tiles
mkdir -p gpurun_out
export JCDF_LIBRARY=$PWD/juliachem.jl_b200/libjcdf_checked.so
python -m pytest tests/test_gpu_block_screened_exchange.py tests/test_gpu_fullsize.py -x -q -m gpu -k "sequence_matches or chain_shapes" > gpurun_out/zzzz_checked_block_screen.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/zzzz_checked_block_screen.log
python tools/sanitizer_cases.py --fuzz 6 block_screened_exchange sliced_exchange_eager > gpurun_out/zzzz_fuzz.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/zzzz_fuzz.log
