#!/usr/bin/env python
"""Full-size consistency of the two kernel families on a synthetic workload the oracle cannot reach: the Fock matrix from
the tile family (ht_mode = 1) and from the per-function family (ht_mode = 2) must agree to 1e-10 per element, and each
must be exactly symmetric.   python tools/check_families.py [workload=gly18_q1167]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import juliachem_jl_b200 as jc  # noqa: E402
from juliachem_jl_b200 import workloads as wl  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "gly18_q1167"
w = wl.WORKLOADS[name]
C, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
mask = wl.screening_mask(w)
out = {}
for mode in (1, 2):
    gd = jc.B200Data(w.N, w.Q, w.occ, 1)
    gd.set_option("ht_mode", mode)
    gd.set_mask(jc.ScreeningData(mask, w.N))
    gd.fill_synthetic_B(w.seed, w.decay_length)
    F, st = gd.fock_build(C, H)
    F, st = gd.fock_build(C, H)
    out[mode] = F.copy()
    li = gd.layout()
    print("ht_mode", mode, "launches", li.kernel_launches_per_build, "finite", bool(np.isfinite(F).all()), "symmetric",
          bool(np.array_equal(F, F.T)), {k: round(v * 1e3, 2) for k, v in st.items() if k in ("W", "K", "V", "J")}, flush=True)
    gd.close()
d = np.abs(out[1] - out[2]).max()
print(name, "N", w.N, "Q", w.Q, "occ", w.occ, "max |F_tiles - F_per_function| =", d, "max |F| =", np.abs(out[1]).max())
sys.exit(0 if d < 1e-10 else 1)
