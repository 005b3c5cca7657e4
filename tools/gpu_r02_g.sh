#!/bin/bash
# round 2, call G: the bounds-checked build (make CHECKED=1) over the smallest case of every kernel family, the parity tests and
# one full-size w50 build — the stand-in for compute-sanitizer memcheck, which is closed on this pool
mkdir -p gpurun_out
export JCDF_LIBRARY=$PWD/juliachem.jl_b200/libjcdf_checked.so
python tools/sanitizer_cases.py --fire > gpurun_out/g_checked_cases.log 2>&1; echo "cases rc=$?"; tail -10 gpurun_out/g_checked_cases.log
python -m pytest tests/test_gpu_parity.py tests/test_gpu_integrals.py -q -x > gpurun_out/g_checked_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/g_checked_pytest.log
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/g_checked_w50.json 2> gpurun_out/g_checked_w50.err; echo "w50 rc=$?"; tail -c 300 gpurun_out/g_checked_w50.json
for wl in c20h42 benzene_dimer gly18_q1167; do python bench.py --workload $wl --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/g_checked_$wl.json 2> gpurun_out/g_checked_$wl.err; echo "$wl rc=$?"; done
grep -l "JCDF_CHECKED violation" gpurun_out/g_checked_* ; echo "violations grep rc=$? (1 = none)"
