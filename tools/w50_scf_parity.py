#!/usr/bin/env python
"""Whole-SCF energy parity of the real w50 (cc-pVDZ / cc-pVDZ-RIFIT, N = 1250, Q = 4800, 500 electrons) with the BENCHED
kernels (default options: per-function half-transform, Coulomb vector from its accumulators, sliced exchange):

  GPU     rhf.run_energy — integrals, Schwarz mask, B, orthogonaliser, DIIS, iteration() all on the device
  oracle  oracle/scf_replay.run_scf (the reference's loop, SCF.jl:399-551) with the oracle's threaded screened Fock build
          (ScreenedDF.jl:227-378, 548-641) on the host cores, fed with the DEVICE-computed S, H, E_nuc, mask and B (read
          back once, 600 rows at a time), so only the per-iteration contraction, DIIS and eigensolver differ.

Writes profiles/r02_w50_scf_parity.json (and a copy under gpurun_out/); exits non-zero unless |E_gpu - E_oracle| <= 1e-8 Eh
with the same number of iterations.   python tools/w50_scf_parity.py [n_molecules=50] [max_iterations=50]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(n_mol=50, niter=50):
    from juliachem_jl_b200 import rhf
    from oracle import df_oracle as orc
    from oracle import scf_replay as sr

    geo = json.load(open(os.path.join(ROOT, "tests", "golden", "w50_geometry.json")))
    mol = rhf.run_basis({"geometry": geo["geometry"][:9 * n_mol], "symbols": geo["symbols"][:3 * n_mol],
                         "molecular_charge": 0}, {"basis": "cc-pVDZ"})
    t0 = time.perf_counter()
    res = rhf.run_energy(mol, {"niter": niter}, num_devices=1, keep_handle=True)
    t_gpu = time.perf_counter() - t0
    gd = res["handle"]
    N, Q, occ, P = res["N"], res["Q"], res["occ"], res["pairs_kept"]
    launches = int(gd.layout().kernel_launches_per_build)
    t0 = time.perf_counter()
    B = np.empty((Q, P), dtype=np.float64, order="F")
    for q0 in range(0, Q, 600):
        q1 = min(Q, q0 + 600)
        B[q0:q1, :] = gd.get_B_rows(0, q0, q1 - q0, P)
    gd.close()
    t_read = time.perf_counter() - t0
    S, H, E_nuc = res["Overlap"], res["Core Hamiltonian"], res["E_nuc"]
    sd = orc.ScreeningData(res["Screen Matrix"])
    cores = len(os.sched_getaffinity(0))
    W_buf = np.zeros((Q, occ, N), dtype=np.float64, order="F")

    def fock(C, it):
        return orc.fock_build_screened_rank_threaded(B, np.asfortranarray(C[:, :occ]), sd, cores, 0, None, W_buf) + H

    t0 = time.perf_counter()
    ref = sr.run_scf(S, H, E_nuc, occ, fock, max_iterations=niter)
    t_cpu = time.perf_counter() - t0
    dE = [abs(a[1] - b[1]) for a, b in zip(res["Iterations"], ref["iterations"])]
    out = {"molecules": n_mol, "N": N, "Q": Q, "occ": occ, "pairs_kept": P, "kernel_launches_per_build": launches,
           "options": "library defaults (the benched kernels)", "E_gpu": res["Energy"], "E_oracle": ref["E"],
           "abs_diff": abs(res["Energy"] - ref["E"]), "tol": 1e-8, "iterations_gpu": len(res["Iterations"]),
           "iterations_oracle": len(ref["iterations"]), "converged_gpu": bool(res["Converged?"]),
           "converged_oracle": bool(ref["converged"]), "max_abs_diff_any_iteration": max(dE) if dE else None,
           "per_iteration": [{"iter": a[0], "E_gpu": a[1], "E_oracle": b[1], "D_rms_gpu": a[3], "D_rms_oracle": b[3]}
                             for a, b in zip(res["Iterations"], ref["iterations"])],
           "seconds": {"gpu_whole_run": round(t_gpu, 2), "read_back_B": round(t_read, 2), "oracle_scf": round(t_cpu, 2)},
           "host_cores": cores}
    out["ok"] = bool(out["abs_diff"] <= 1e-8 and out["iterations_gpu"] == out["iterations_oracle"] and out["converged_gpu"]
                     and out["converged_oracle"])
    return out


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    it = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    out = run(n, it)
    for d in ("profiles", "gpurun_out"):
        os.makedirs(os.path.join(ROOT, d), exist_ok=True)
        json.dump(out, open(os.path.join(ROOT, d, "r02_w50_scf_parity%s.json" % ("" if n == 50 else "_%d" % n)), "w"), indent=1)
    print(json.dumps({k: v for k, v in out.items() if k != "per_iteration"}))
    sys.exit(0 if out["ok"] else 1)
