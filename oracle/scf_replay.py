"""Replay of the reference's DF-RHF SCF driver around the Fock build (TEST INFRASTRUCTURE — never shipped).

Restates, in numpy, the part of src/rhf/energy/SCF.jl that surrounds `df_rhf_fock_build!`, so a Fock-build
implementation (the oracle's, or libjcdf through its C-ABI in the GPU tests) can be driven through the very iteration
sequence the reference logged in water_ccpvdz_out.log:

  rhf_kernel         SCF.jl:69-262   H = T + V; X = S^-1/2 from the eigen-decomposition of S (:141-163, eigenvalues
                                     below 1e-6 dropped); hcore guess = iteration(H) (:178-181)
  scf_cycles_kernel  SCF.jl:340-592  per iteration: F = df_rhf_fock_build!(C) (:463); DIIS on e = FDS - (FDS)^T
                                     (:471-501, newest first, B_dim capped at ndiis, DIIS() EnergyHelpers.jl:234-258);
                                     damping x = 1/log_50(50 dE) when dE >= 1 (:504-505); iteration() (:512);
                                     dE, D_rms = ||D - D_old||_F (:521-525); DF convergence test (:527-530)
  iteration          SCF.jl:1072-1125 (oracle/df_oracle.py: iteration)

`fock_build(C, iteration)` must return the full Fock matrix H + 2J - K for the occupied block C[:, :nocc].
"""
from __future__ import annotations

import math

import numpy as np
import scipy.linalg as sla

from . import df_oracle as orc


def orthogonalizer(S: np.ndarray, threshold: float = 1.0e-6) -> np.ndarray:
    """SCF.jl:141-163: X = U_good diag(s_good^-1/2) U_good^T."""
    w, U = sla.eigh(S)
    good = w >= threshold
    return (U[:, good] * (w[good] ** -0.5)) @ U[:, good].T


def diis_extrapolate(e_array, F_array, B_dim: int) -> np.ndarray:
    """DIIS() (EnergyHelpers.jl:234-258): B[i,j] = <e_i, e_j>, border -1, rhs (0..0,-1), sysv!('U')."""
    B = np.zeros((B_dim + 1, B_dim + 1))
    for i in range(B_dim):
        for j in range(B_dim):
            B[i, j] = np.vdot(e_array[i], e_array[j])
        B[i, B_dim] = B[B_dim, i] = -1.0
    rhs = np.zeros(B_dim + 1)
    rhs[B_dim] = -1.0
    coeff = sla.solve(B, rhs, assume_a="sym")
    F = np.zeros_like(F_array[0])
    for i in range(B_dim):
        F += coeff[i] * F_array[i]
    return F


def run_scf(S, Hcore, E_nuc: float, nocc: int, fock_build, ndiis: int = 10, dele: float = 1e-6, rmsd: float = 1e-6,
            max_iterations: int = 50, record=None, iteration_fn=None, pass_previous_fock: bool = False):
    """Returns dict(E, converged, iterations=[(iter, E, dE, D_rms)], C, D, F).  iteration_fn(F, H, X, nocc) ->
    (E_elec, evals, C, D) replaces the oracle's iteration() (e.g. with libjcdf's on-device jcdf_eigh in GPU tests).
    pass_previous_fock: call fock_build(C, iteration, two_electron_fock) where the third argument is what
    scf_data.two_electron_fock holds at that moment — None at iteration 1 (zeros, DensityFitting.jl:45), afterwards the
    matrix the previous call returned AFTER this loop's DIIS and damping, which act on that very array in place
    (SCF.jl:463, 491, 505: `F` is scf_data.two_electron_fock).  Only df_exchange_screen reads it."""
    N = S.shape[0]
    ortho = orthogonalizer(S)
    iteration = iteration_fn or orc.iteration
    # hcore guess: iteration(F = H) (SCF.jl:178-181)
    E_elec, evals, C, D = iteration(np.array(Hcore), Hcore, ortho, nocc)
    F_old = np.array(Hcore)
    E_old, dE, D_rms = 0.0, 1.0, 1.0
    e_array, F_array = [], []
    B_dim = 1
    log = []
    it = 1
    converged = False
    F = np.array(Hcore)
    while True:
        if pass_previous_fock:
            F = np.array(fock_build(C, it, None if it == 1 else F), dtype=np.float64, order="F")
        else:
            F = np.array(fock_build(C, it), dtype=np.float64, order="F")
        if record is not None:
            record(it, "F", F.copy())
        if ndiis > 0:
            FDS = F @ D @ S
            err = FDS - FDS.T
            e_array = [err] + e_array[:ndiis - 1]
            F_array = [F.copy()] + F_array[:ndiis - 1]
            if it > 1:
                B_dim = min(B_dim + 1, ndiis)
                try:
                    F = diis_extrapolate(e_array, F_array, B_dim)
                except Exception:  # "Faulty DIIS!" (SCF.jl:493-499)
                    B_dim = 2
        x = 1.0 / math.log(50.0 * dE, 50.0) if dE >= 1.0 else 1.0
        F = (1.0 - x) * F_old + x * F
        F_old = F.copy()
        D_old = D
        E_elec, evals, C, D = iteration(F, Hcore, ortho, nocc)
        D_rms = float(np.sqrt(np.vdot(D - D_old, D - D_old)))
        E = E_elec + E_nuc
        dE = E - E_old
        log.append((it, E, dE, D_rms))
        if abs(dE) <= dele and D_rms <= rmsd:
            converged = True
            break
        if it >= max_iterations:
            break
        it += 1
        E_old = E
    return {"E": E, "converged": converged, "iterations": log, "C": C, "D": D, "F": F, "evals": evals}
