#!/usr/bin/env python
"""Spread of the per-CTA busy time of the exchange kernel on the real w50 (cost-balanced work lists vs round-robin).
   python tools/exchange_balance.py [workload=w50]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import juliachem_jl_b200 as jc  # noqa: E402
from juliachem_jl_b200 import rhf, workloads as wl  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "w50"
w = wl.WORKLOADS[name]
C, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
for balance in (1, 0):
    gd = jc.B200Data(w.N, w.Q, w.occ, 1)
    gd.set_option("exchange_balance", balance)
    if w.real_molecules:
        mol = rhf.run_basis(wl.real_molecule(w), {"basis": "cc-pVDZ"})
        gd.set_basis(0, *mol["primary"].args())
        gd.set_basis(1, *mol["auxiliary"].args())
        gd.compute_two_center(False)
        gd.compute_schwarz_mask(1e-5)
        gd.compute_B()
    else:
        gd.set_mask(jc.ScreeningData(wl.screening_mask(w), w.N))
        gd.fill_synthetic_B(w.seed, w.decay_length)
    for _ in range(3):
        F, st = gd.fock_build(C, H)
    cyc = gd.exchange_cta_cycles().astype(np.float64)
    print("balance", balance, "K ms", round(st["K"] * 1e3, 3), "CTAs", len(cyc), "cycles min/mean/max (M)",
          round(cyc.min() / 1e6, 2), round(cyc.mean() / 1e6, 2), round(cyc.max() / 1e6, 2),
          "max/mean", round(cyc.max() / cyc.mean(), 4), "sorted tail", np.sort(cyc)[-5:] / 1e6, flush=True)
    gd.close()
