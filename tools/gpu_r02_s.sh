#!/bin/bash
# round 2, call S: the checked build (bounds checks + schedule fuzzing) on the kernels added since call L — two CTAs per SM,
# block-screened epilogue — then every family again; margins of the block-screened SCF test
mkdir -p gpurun_out
( export JCDF_LIBRARY=$PWD/juliachem.jl_b200/libjcdf_checked.so; python tools/sanitizer_cases.py --fuzz 8 ) > gpurun_out/s_fuzz_cases.log 2>&1; echo "fuzz cases rc=$?"; tail -16 gpurun_out/s_fuzz_cases.log
( export JCDF_LIBRARY=$PWD/juliachem.jl_b200/libjcdf_checked.so; python -m pytest tests/test_gpu_block_screened_exchange.py tests/test_gpu_parity.py -x -q -m gpu -k "block_screened or warp_layouts or dropped or seam" ) > gpurun_out/s_checked_pytest.log 2>&1; echo "checked pytest rc=$?"; tail -4 gpurun_out/s_checked_pytest.log
python -m pytest tests/test_gpu_block_screened_exchange.py -x -q -m gpu -s -k resident > gpurun_out/s_scf_margins.log 2>&1; echo "scf rc=$?"; grep "block-screened SCF" gpurun_out/s_scf_margins.log
