#!/bin/bash
# round 2, call L: schedule fuzzing with the checked build (stand-in for racecheck): every family, 12 seeds; one full-size w50
# build under fuzzing against the fixture
mkdir -p gpurun_out
export JCDF_LIBRARY=$PWD/juliachem.jl_b200/libjcdf_checked.so
python tools/sanitizer_cases.py --fuzz 12 > gpurun_out/l_fuzz_cases.log 2>&1; echo "fuzz cases rc=$?"; tail -12 gpurun_out/l_fuzz_cases.log
python - > gpurun_out/l_fuzz_w50.log 2>&1 <<'PY'
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
from test_gpu_fullsize import _w50
from juliachem_jl_b200 import workloads as wl
w = wl.WORKLOADS["w50"]
C, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
_, gd = _w50(0)
F0 = gd.fock_build(C, H)[0].copy()
fx = json.load(open("tests/golden/w50_fock_fixture.json")); idx = np.array(fx["index"])
print("seed 0: fixture err %.2e" % np.abs(F0[idx[:, 0], idx[:, 1]] - np.array(fx["value"])).max(), flush=True)
for seed in (1, 2, 3):
    gd.checked_set_fuzz(seed)
    t0 = time.perf_counter(); F = gd.fock_build(C, None)[0]; dt = time.perf_counter() - t0
    print("seed %d: identical=%s max|dF|=%.1e  build %.1f ms (sleeps included)" % (seed, np.array_equal(F, F0), np.abs(F - F0).max(), dt * 1e3), flush=True)
    assert np.array_equal(F, F0)
gd.close()
PY
echo "fuzz w50 rc=$?"; cat gpurun_out/l_fuzz_w50.log | tail -5
