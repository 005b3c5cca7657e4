"""Builds the water cc-pVDZ / cc-pVDZ-RIFIT problem of the golden fixture tests/golden/water_ccpvdz_rifit.json with the
oracle's own integrals (TEST INFRASTRUCTURE).  Used by tests/ to pin oracle/df_oracle.py — and through it libjcdf —
against the reference's logged run."""
from __future__ import annotations

import json
import os

import numpy as np

from . import df_oracle as orc
from . import gaussian_integrals as gi

GOLDEN = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                      "water_ccpvdz_rifit.json")


def load_golden(path: str = GOLDEN) -> dict:
    return json.load(open(path))


def shells_from(golden_basis, geometry):
    shells = []
    for ia, atom in enumerate(golden_basis):
        c = np.array(geometry[ia]["xyz_bohr"], dtype=np.float64)
        for s in atom["shells"]:
            shells.append(gi.Shell(int(s["l"]), np.array(s["exps"]), np.array(s["coefs"]), c, ia))
    return shells


def build_problem(g: dict | None = None) -> dict:
    """S, T, V, H, E_nuc, (P|Q) [lower triangle valid], (P|mn) as T3[Q, N, N], B = L^-1 T (dense, (Q,N,N) 'F')."""
    g = g or load_golden()
    prim = shells_from(g["basis"], g["geometry"])
    aux = shells_from(g["auxiliary_basis"], g["geometry"])
    centers = [a["xyz_bohr"] for a in g["geometry"]]
    S, T = gi.overlap_kinetic(prim)
    V = gi.nuclear_attraction(prim, g["charges"], centers)
    J = gi.two_center_integrals(aux)
    T3 = gi.three_center_integrals(aux, prim)
    Q, N = T3.shape[0], T3.shape[1]
    T2 = np.asfortranarray(T3.reshape(Q, N * N))  # column (mu,nu) -> mu*N+nu == nu + N*mu by symmetry
    B = orc.calculate_B(np.tril(J), T2).reshape((Q, N, N), order="F")
    return {"S": S, "T": T, "V": V, "H": np.asfortranarray(T + V), "E_nuc": gi.nuclear_repulsion(g["charges"], centers),
            "J_PQ": J, "T3": T3, "B": np.asfortranarray(B), "nocc": g["settings"]["n_electrons"] // 2,
            "aux_shell_nbas": [s.nbas for s in aux], "N": N, "Q": Q}
