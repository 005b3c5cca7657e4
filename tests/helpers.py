"""Shared builders for the parity tests: seeded inputs -> (oracle result, arrays for the C-ABI)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import df_oracle as orc  # noqa: E402  (tests are allowed to import the oracle)
import juliachem_jl_b200 as jc  # noqa: E402
from juliachem_jl_b200 import workloads as wl  # noqa: E402


def random_inputs(N, Q, occ, seed, mask=None, decay=0.0):
    """Symmetric-in-(p,q) B (Q,N,N) 'F', zero outside mask; orthonormal C_occ; symmetric H."""
    rng = np.random.default_rng(seed)
    B = rng.standard_normal((Q, N, N)) / np.sqrt(Q)
    B = 0.5 * (B + B.transpose(0, 2, 1))
    if decay > 0:
        idx = np.arange(N)
        B *= np.exp(-np.abs(idx[:, None] - idx[None, :]) / decay)[None]
    if mask is not None:
        B *= mask[None]
    B = np.asfortranarray(B)
    A = rng.standard_normal((N, occ))
    C_occ = np.asfortranarray(np.linalg.qr(A)[0][:, :occ])
    Hs = rng.standard_normal((N, N))
    H = np.asfortranarray(0.5 * (Hs + Hs.T))
    return B, C_occ, H


def banded_mask(N, half_width, seed=0, speckle=True):
    idx = np.arange(N)
    d = np.abs(idx[:, None] - idx[None, :])
    keep = d <= half_width
    if speckle:
        rng = np.random.default_rng(seed)
        r = rng.random((N, N))
        r = np.minimum(r, r.T)
        keep &= ~((d > 0.8 * half_width) & (r < 0.3))
    np.fill_diagonal(keep, True)
    return np.asfortranarray(keep)


def oracle_fock(B_dense, C_occ, H, mask, shards=None):
    """Reference-algorithm Fock: screened algorithm when a mask is given, dense algorithm otherwise; shards = list of
    ranges emulating ranks/devices."""
    Q = B_dense.shape[0]
    N = B_dense.shape[1]
    occ = C_occ.shape[1]
    shards = shards or [range(0, Q)]
    Cfull = np.zeros((N, N), order="F")
    Cfull[:, :occ] = C_occ
    if mask is None:
        Bs = [np.asfortranarray(B_dense[r.start:r.stop]) for r in shards]
        return orc.df_rhf_fock_build(Bs, Cfull, occ, H, None, "dense")
    sd = orc.ScreeningData(mask)
    Bs = [orc.dense_to_screened(np.asfortranarray(B_dense[r.start:r.stop]), sd) for r in shards]
    return orc.df_rhf_fock_build(Bs, Cfull, occ, H, sd, "screened")


def reference_layout_rows(B_dense, q_range, mask):
    """(len(q_range), P) column-major in the reference pair order, as jcdf_set_B / jcdf_upload_three_center expect."""
    Bq = np.asfortranarray(B_dense[q_range.start:q_range.stop])
    Qr, N, _ = Bq.shape
    B2 = Bq.reshape((Qr, N * N), order="F")
    if mask is None:
        return np.asfortranarray(B2)
    return np.asfortranarray(B2[:, np.asfortranarray(mask).ravel(order="F")])
