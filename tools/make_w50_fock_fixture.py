#!/usr/bin/env python
"""Makes tests/golden/w50_fock_fixture.json on a B200 box: the ORACLE's Fock matrix of the real w50 (cc-pVDZ /
cc-pVDZ-RIFIT, N = 1250, Q = 4800, occ = 250) for bench.py's seeded C_occ and H, sampled at 4096 seeded positions, plus
its Frobenius norm.  B = L^-1 (P|mn) is computed on the device, read back in the reference layout 1200 rows at a time
(jcdf_get_B_rows) and contracted by oracle.fock_build_screened_slabs (the reference's threaded screened algorithm,
ScreenedDF.jl:227-378, 548-641) on the host cores.  The GPU Fock matrix of the default kernels is compared first
(< 1e-10 per element) and the measured error is recorded.  bench.py asserts these samples after every timed run at
every GPU count; tests/test_gpu_fullsize.py re-derives them.

    python tools/make_w50_fock_fixture.py [out.json]      (also copied to gpurun_out/ so it travels back)
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    from test_gpu_fullsize import _w50, oracle_fock_from_device_B  # the same builders the parity test uses
    from juliachem_jl_b200 import workloads as wl
    w = wl.WORKLOADS["w50"]
    C, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
    _, gd = _w50(0)
    F = gd.fock_build(C, H)[0].copy()
    t0 = time.perf_counter()
    F_orc = oracle_fock_from_device_B(gd, w.Q, C, H, wl.real_mask(w), 1200)
    t_orc = time.perf_counter() - t0
    gd.close()
    err = float(np.abs(F - F_orc).max())
    print("max|F_gpu - F_oracle| = %.3e, max|F| = %.4f, oracle + read-back %.1f s" % (err, np.abs(F_orc).max(), t_orc))
    if not err < 1e-10:
        raise SystemExit("GPU and oracle disagree: no fixture written")
    rng = np.random.default_rng(20261105)
    p = rng.integers(0, w.N, size=4096)
    q = rng.integers(0, w.N, size=4096)
    p[:64] = q[:64]  # some diagonal elements
    out = {"workload": "w50", "N": w.N, "Q": w.Q, "occ": w.occ, "seeds": {"C_occ": 2000 + w.seed, "H": 3000 + w.seed},
           "made_by": "tools/make_w50_fock_fixture.py", "values_from": "oracle (fock_build_screened_slabs, 4 slabs of 1200 rows)",
           "gpu_vs_oracle_max_abs_err_when_made": err, "max_abs_F": float(np.abs(F_orc).max()),
           "frobenius": float(np.linalg.norm(F_orc)), "trace": float(np.trace(F_orc)),
           "index": np.stack([p, q], axis=1).tolist(), "value": [float(x) for x in F_orc[p, q]]}
    path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "w50_fock_fixture.json")
    json.dump(out, open(path, "w"))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "w50_fock_fixture.json"), "w"))
    print("wrote", path)


if __name__ == "__main__":
    main()
