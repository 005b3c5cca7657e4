#!/bin/bash
# round 2, call J: the reference's own S22 acceptance suite (H/C/O complexes) on the device, as records and as tests
mkdir -p gpurun_out
python tools/s22_suite.py 1 gpurun_out/r02_s22_suite.json > gpurun_out/j_s22.log 2>&1; echo "suite rc=$?"; cut -c1-330 gpurun_out/j_s22.log | tail -12
python -m pytest tests/test_gpu_s22_reference_suite.py -q -x > gpurun_out/j_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/j_pytest.log
