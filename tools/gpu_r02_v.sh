#!/bin/bash
# round 2, call V (2 GPUs): the sequence that failed in call U — multi-device tests (libjcdf maps NCCL), then the test that imports
# torch afterwards, in ONE process — and the rest of the multi-device cases
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_block_screened_exchange.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q -m gpu -k "several_devices or multi_device or no_alloc or devices" ) > gpurun_out/v_nccl_then_torch.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/v_nccl_then_torch.log
python - <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import juliachem_jl_b200 as jc
print("JCDF_NCCL_LIBRARY =", os.environ.get("JCDF_NCCL_LIBRARY"))
PY
