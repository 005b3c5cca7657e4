"""Synthetic inputs of the shapes BASELINE.json's configs name: sizes (N, Q, occ), a geometry-driven
symmetric screening mask with a stated survival fraction, an orthonormal C_occ, a symmetric H, and the host twin of
libjcdf's on-device synthetic B generator (jcdf_fill_synthetic_B) so tests can rebuild any B row bit for bit.

Real integrals are not available for these molecules in this image (no basis-set database), so the
benchmark tensors are synthetic; shapes, sparsity structure and arithmetic are those of the named configuration."""
from __future__ import annotations

from dataclasses import dataclass

import math

import numpy as np


@dataclass(frozen=True)
class Workload:
    name: str
    config: str          # the BASELINE.json config string it stands for
    N: int
    Q: int
    occ: int
    geometry: str        # "none" | "chain" | "cluster"
    atoms_heavy: int     # heavy atoms / molecules used to lay functions out in space
    cutoff: float        # kept iff centre distance <= cutoff (in units of the nearest-neighbour spacing); 0 = keep all
    decay_length: float  # index-distance decay of the synthetic B values (0 = none)
    seed: int
    real_molecules: int = 0  # > 0: the first n molecules of tests/golden/w50_geometry.json with REAL cc-pVDZ/cc-pVDZ-RIFIT
                             # integrals computed on the device (mask from the reference's Schwarz rule, df_sigma = 1e-5)
    real_s22: int = 0        # > 0: that S22 complex (tests/golden/s22_hco.json, the reference's example_inputs/S22) with REAL
                             # 6-31G(2df,p) / cc-pVTZ-JKFIT integrals computed on the device
    basis: str = "cc-pVDZ"   # basis-set name of a real workload (rhf.load_basis_library)


WORKLOADS = {
    "water": Workload("water", "water DF-RHF cc-pVDZ / cc-pVDZ-RI single SCF", 25, 96, 5, "none", 1, 0.0, 0.0, 1001),
    "benzene_dimer": Workload("benzene_dimer", "S22 benzene dimer DF-RHF cc-pVTZ / cc-pVTZ-RI on 1 B200", 600, 1632, 42,
                              "none", 12, 0.0, 0.0, 1002),
    # the configuration BASELINE.json's metric is quoted on, with its real integrals and real Schwarz mask
    "w50": Workload("w50", "w50 water cluster DF-RHF cc-pVDZ Q-sharded at 1/2/4/8 B200", 1250, 4800, 250, "real", 50,
                    0.0, 0.0, 1003, real_molecules=50),
    # same shape, synthetic B with no screening at all (upper bound of the work: every pair kept)
    "w50_dense": Workload("w50_dense", "w50 water cluster DF-RHF cc-pVDZ shape, synthetic unscreened B", 1250, 4800, 250,
                          "none", 50, 0.0, 400.0, 1003),
    "w50_screened": Workload("w50_screened", "w50 water cluster DF-RHF cc-pVDZ, Schwarz mask", 1250, 4800, 250,
                             "cluster", 50, 2.6, 400.0, 1013),
    # one rank's share of w50 when Q is sharded over 8 GPUs (profiling stand-in for the 8-GPU per-device work)
    "w50_q600": Workload("w50_q600", "w50 water cluster DF-RHF cc-pVDZ, the 600 auxiliary rows one of 8 GPUs owns", 1250, 600,
                         250, "none", 50, 0.0, 400.0, 1003),
    # the S22 benzene dimer (parallel displaced) itself with real integrals and its real Schwarz mask (67 % of the pairs
    # survive) — in the basis the reference's own S22 inputs use, because cc-pVTZ data are not in the snapshot
    "benzene_dimer_real": Workload("benzene_dimer_real", "S22 benzene dimer DF-RHF, real 6-31G(2df,p) / cc-pVTZ-JKFIT integrals on 1 B200",
                                   468, 1572, 42, "real", 24, 0.0, 0.0, 1012, real_molecules=1, real_s22=11, basis="6-31G(2df,p)"),
    "c20h42": Workload("c20h42", "C20H42 alkane DF-RHF cc-pVTZ with shell-pair screening mask", 1330, 3490, 81,
                       "chain", 20, 6.5, 300.0, 1004),
    "gly18": Workload("gly18", "glycine oligomer DF-RHF cc-pVTZ, largest B tensor, 8 B200", 3395, 9333, 275, "chain",
                      90, 14.0, 600.0, 1005),
    # one rank's share of gly18 when Q is sharded over 8 GPUs (profiling stand-in for the 8-GPU per-device work)
    "gly18_q1167": Workload("gly18_q1167", "glycine oligomer DF-RHF cc-pVTZ, the 1167 auxiliary rows one of 8 GPUs owns",
                            3395, 1167, 275, "chain", 90, 14.0, 600.0, 1005),
    # reduced shapes for parity tests the oracle finishes in seconds
    "tiny": Workload("tiny", "parity-test shape", 70, 40, 9, "chain", 7, 2.5, 20.0, 1100),
    "small": Workload("small", "parity-test shape", 203, 131, 27, "cluster", 12, 1.9, 60.0, 1101),
}


def function_centres(w: Workload) -> np.ndarray:
    """(N, 3) coordinates: functions are dealt to `atoms_heavy` sites in index order (like shells sorted by atom);
    chain = sites on a line at unit spacing (alkane / peptide backbone order of the input files);
    cluster = jittered cubic lattice at unit spacing (water cluster)."""
    rng = np.random.default_rng(w.seed + 7)
    ns = max(1, w.atoms_heavy)
    if w.geometry == "chain":
        sites = np.stack([np.arange(ns, dtype=np.float64), np.zeros(ns), np.zeros(ns)], axis=1)
        sites += rng.normal(scale=0.08, size=sites.shape)
    elif w.geometry == "cluster":
        side = int(np.ceil(ns ** (1.0 / 3.0)))
        grid = np.array([(i, j, k) for i in range(side) for j in range(side) for k in range(side)], dtype=np.float64)
        sites = grid[:ns] + rng.normal(scale=0.15, size=(ns, 3))
    else:
        sites = np.zeros((ns, 3))
    owner = (np.arange(w.N, dtype=np.int64) * ns) // w.N
    return sites[owner]


def screening_mask(w: Workload):
    """Symmetric Bool (N, N) column-major mask with the diagonal kept, or None for unscreened workloads.
    Kept iff the two functions' sites are within `cutoff`; pairs in the outer 15 % shell of the cutoff are thinned by
    a symmetric hash (the speckle at the band edge seen in the reference's C20H42_screening_matrix.png)."""
    if w.real_molecules:
        return real_mask(w)
    if w.geometry == "none" or w.cutoff <= 0:
        return None
    x = function_centres(w)
    d = np.sqrt(((x[:, None, :] - x[None, :, :]) ** 2).sum(axis=2))
    keep = d <= w.cutoff
    edge = (d > 0.85 * w.cutoff) & keep
    p, q = np.indices((w.N, w.N))
    lo, hi = np.minimum(p, q).astype(np.uint64), np.maximum(p, q).astype(np.uint64)
    h = splitmix64((lo << np.uint64(21)) | hi ^ np.uint64(w.seed))
    keep &= ~(edge & ((h & np.uint64(3)) == 0))
    np.fill_diagonal(keep, True)
    return np.asfortranarray(keep)


def real_molecule(w: Workload) -> dict:
    """JSON-contract molecule object of a real workload (first n waters of the reference's w50 input)."""
    import json
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if w.real_s22:
        m = json.load(open(os.path.join(root, "tests", "golden", "s22_hco.json")))["molecules"][str(w.real_s22)]
        return {"geometry": m["geometry"], "symbols": m["symbols"], "molecular_charge": m["molecular_charge"]}
    geo = json.load(open(os.path.join(root, "tests", "golden", "w50_geometry.json")))
    n = w.real_molecules
    return {"geometry": geo["geometry"][:9 * n], "symbols": geo["symbols"][:3 * n], "molecular_charge": 0}


def real_mask(w: Workload):
    """The Schwarz mask libjcdf computes for the real w50 (df_sigma = 1e-5), committed as a fixture so the CPU legs of
    bench.py can run without a GPU (tests/golden/w50_mask_sigma1e-5.npy, made by tools/dump_w50_mask.py)."""
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if w.real_s22:
        bits = np.load(os.path.join(root, "tests", "golden", "s22_%d_mask_sigma1e-5.npy" % w.real_s22))
    elif w.real_molecules != 50:
        raise ValueError("only the 50-molecule mask is committed")
    else:
        bits = np.load(os.path.join(root, "tests", "golden", "w50_mask_sigma1e-5.npy"))
    m = np.unpackbits(bits)[:w.N * w.N].reshape(w.N, w.N).astype(bool)
    return np.asfortranarray(m)


def occupied_coefficients(w: Workload, seed_offset: int = 0) -> np.ndarray:
    """First occ columns of the Q factor of a seeded Gaussian matrix: orthonormal columns."""
    rng = np.random.default_rng(2000 + w.seed + seed_offset)
    A = rng.standard_normal((w.N, w.occ))
    Qf, _ = np.linalg.qr(A)
    return np.asfortranarray(Qf[:, :w.occ])


def core_hamiltonian(w: Workload) -> np.ndarray:
    rng = np.random.default_rng(3000 + w.seed)
    A = rng.standard_normal((w.N, w.N))
    return np.asfortranarray(0.5 * (A + A.T))


# ------------------------------------------------------------------------------------------------------------------
# host twin of csrc/kernels_misc.cu: splitmix64 / synth_isum / jcdf_synthetic_value
# ------------------------------------------------------------------------------------------------------------------
def splitmix64(x: np.ndarray) -> np.ndarray:
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = x + np.uint64(0x9E3779B97F4A7C15)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def synthetic_B_rows(w: Workload, q_range: range, mask) -> np.ndarray:
    """Dense (len(q_range), N, N) column-major rows of the synthetic B, bit-identical to jcdf_fill_synthetic_B(seed=
    w.seed, decay_length=w.decay_length): value = isum(seed, Q, lo, hi) * cdec[hi - lo], zero outside the mask."""
    N = w.N
    p, q = np.indices((N, N))
    lo, hi = np.minimum(p, q).astype(np.uint64), np.maximum(p, q).astype(np.uint64)
    base = (lo << np.uint64(21)) | hi
    sigma = 65536.0 / 1.7320508075688772
    dist = np.arange(N, dtype=np.float64)
    # math.exp is libm's exp, the function synth_table (kernels_misc.cu) calls; numpy's SIMD exp can differ in the last bit
    dec = np.array([math.exp(-d / w.decay_length) for d in dist]) if w.decay_length > 0 else np.ones(N)
    cdec = dec / (sigma * np.sqrt(float(w.Q)))
    scale = cdec[(hi - lo).astype(np.int64)]
    if mask is not None:
        scale = np.where(mask, scale, 0.0)
    sseed = splitmix64(np.uint64(w.seed))
    out = np.empty((len(q_range), N, N), dtype=np.float64, order="F")
    for k, qa in enumerate(q_range):
        h = splitmix64((base | (np.uint64(qa) << np.uint64(42))) ^ sseed)
        s = ((h & np.uint64(0xFFFF)).astype(np.int64) + ((h >> np.uint64(16)) & np.uint64(0xFFFF)).astype(np.int64)
             + ((h >> np.uint64(32)) & np.uint64(0xFFFF)).astype(np.int64)
             + ((h >> np.uint64(48)) & np.uint64(0xFFFF)).astype(np.int64)) - 2 * 65535
        out[k] = s.astype(np.float64) * scale
    return out


def exchange_dropped_elements(mask, n_blocks: int):
    """Which elements of F the block-screened exchange leaves uncomputed (option "exchange_block_screen", the reference's
    df_exchange_screen: calculate_exchange_block_screen_matrix, ScreenedDF.jl:385-457) as a Bool (N, N) matrix in the caller's
    order, plus (lower-triangular blocks, blocks kept).  Below 100 functions one block; width N // n_blocks; a block is dropped
    iff its sub-block of the mask holds no pair; the remainder strip is never screened."""
    if mask is None:
        raise ValueError("the dense statement of the contraction has no K blocks (NULL mask)")
    N = mask.shape[0]
    if n_blocks <= 0:
        return np.zeros((N, N), dtype=bool), 0, 0
    width = N if N < 100 else N // n_blocks
    nb = 1 if N < 100 else n_blocks
    sub = np.asarray(mask)[:nb * width, :nb * width].reshape(nb, width, nb, width).any(axis=(1, 3))
    out = np.zeros((N, N), dtype=bool)
    out[:nb * width, :nb * width] = ~np.kron(sub, np.ones((width, width), dtype=bool))
    return out, nb * (nb + 1) // 2, int(np.tril(sub).sum())


def algorithmic_work(N: int, Q: int, occ: int, pairs_kept: int, pairs_tri: int) -> dict:
    """SURVEY.md §8(d) per-build figures for a shard of Q auxiliary functions."""
    return {
        "j_bytes": 8.0 * Q * pairs_tri + 16.0 * N * N,
        "j_flops": 4.0 * Q * pairs_tri,
        "w_flops": 2.0 * Q * occ * pairs_kept,
        "k_flops": 1.0 * Q * occ * N * (N + 1),
        "dense_equiv_flops": 4.0 * Q * occ * N * N,
    }
