#!/bin/bash
# round 2, call K: the real S22 benzene dimer as a bench workload (with its CPU leg and in-run oracle parity), S22 tests
mkdir -p gpurun_out
python bench.py --workload benzene_dimer_real --steps 20 --warmup 3 > gpurun_out/k_benzene_real.json 2> gpurun_out/k_benzene_real.err; echo "rc=$?"; tail -c 1300 gpurun_out/k_benzene_real.json; tail -2 gpurun_out/k_benzene_real.err
python bench.py --impl reference --workload benzene_dimer_real --steps 10 --warmup 1 > gpurun_out/k_benzene_real_reference.json 2>&1; head -c 300 gpurun_out/k_benzene_real_reference.json
python -m pytest tests/test_gpu_s22_reference_suite.py -q > gpurun_out/k_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/k_pytest.log
