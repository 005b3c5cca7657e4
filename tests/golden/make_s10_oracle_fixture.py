#!/usr/bin/env python
"""Oracle results for S22 system 10 (benzene-methane, 6-31G(2df,p) / cc-pVTZ-JKFIT, N = 297, Naux = 1022, real Schwarz
mask at df_sigma = 1e-5: 76 % of the pairs survive) as a compact fixture for the GPU tests, which cannot afford the
oracle's integrals (about 40 s on 8 cores; tests/test_oracle_s10_golden.py re-derives and checks this file on the CPU):

    tests/golden/s10_benzene_methane_oracle.npz
        mask_bits   packed basis_function_screen_matrix (reference rule, SchwarzScreening.jl:9-83)
        H           core Hamiltonian T + V
        C_occ       occupied block of the hcore guess (iteration(F = H), SCF.jl:178-181)
        F1          Fock matrix of SCF iteration 1 from the reference's DEFAULT screened algorithm on 3 ranks sharded by
                    auxiliary shells (DynamicLoad.jl:160-203) — oracle.df_rhf_fock_build
        E_nuc, E1   nuclear repulsion; total energy of iteration 1 (the reference logged -172.5375557595)

Run where the oracle can run (any CPU box):   python tests/golden/make_s10_oracle_fixture.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import df_oracle as orc  # noqa: E402
from oracle import molecule_case as mc  # noqa: E402
from oracle import scf_replay as sr  # noqa: E402


def compute(workers=None):
    g = json.load(open(os.path.join(HERE, "s10_631g2dfp_jkfit.json")))
    p = mc.build_problem(g["benzene_methane"], g["settings"]["df_sigma"], g["n_ranks"], workers or os.cpu_count() or 1)
    H, nocc, sd = p["H"], p["nocc"], p["sd"]
    X = sr.orthogonalizer(p["S"])
    E0, _, C, _ = orc.iteration(np.array(H), H, X, nocc)
    F1 = orc.df_rhf_fock_build(p["B_shards"], C, nocc, H, sd, "screened")
    E1, ev, _, _ = orc.iteration(F1, H, X, nocc)
    return {"mask_bits": np.packbits(np.asarray(p["fn_mask"], dtype=np.uint8).ravel(order="F")), "H": H,
            "C_occ": np.asfortranarray(C[:, :nocc]), "F1": F1, "E_nuc": np.float64(p["E_nuc"]),
            "E_guess": np.float64(E0 + p["E_nuc"]), "E1": np.float64(E1 + p["E_nuc"]),
            "homo_lumo_iter1": np.array([ev[nocc - 1], ev[nocc]]), "shard_sizes": np.array([len(r) for r in p["ranges"]])}


if __name__ == "__main__":
    out = compute()
    path = os.path.join(HERE, "s10_benzene_methane_oracle.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; E1 =", float(out["E1"]), "shards", out["shard_sizes"])
