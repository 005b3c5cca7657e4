"""Oracle parity at the sizes the benchmark is quoted on.

Real w50 (cc-pVDZ / cc-pVDZ-RIFIT, N = 1250, Q = 4800, occ = 250; integrals, Schwarz mask and B computed on the device):
B is pulled back through jcdf_get_B_rows in the reference's (Q x P) layout, 1200 rows at a time, and the oracle's
restatement of the reference's threaded screened algorithm (ScreenedDF.jl:227-378, 548-641) builds the same Fock matrix
on the host cores — each slab is one emulated MPI rank, the partials are summed as DensityFitting.jl:68-71 does.  The
DEFAULT option set (per-function half-transform, Coulomb vector from its accumulators, sliced exchange) must agree to
1e-10 per element; so must the tile family, two ragged Q shards, and the committed fixture every bench.py run asserts
(tests/golden/w50_fock_fixture.json).  10 exchange tile rows, 38 chunks of 128 auxiliary rows, 78 contraction slices.

The other BASELINE shapes (benzene dimer, C20H42, one gly18 shard; synthetic B) are checked the same way on a
slab of auxiliary rows — the build is a plain sum over rows, and a slab keeps the oracle under a minute."""
import json
import os

import numpy as np
import pytest

from helpers import ROOT, jc, orc, wl

pytestmark = pytest.mark.gpu
TOL = 1e-10
CORES = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def _w50(ht_mode, shard=None):
    from juliachem_jl_b200 import rhf
    w = wl.WORKLOADS["w50"]
    gd = jc.B200Data(w.N, w.Q, w.occ, 1)
    if shard is not None:
        gd.set_shards([shard])
    gd.set_option("ht_mode", ht_mode)
    mol = rhf.run_basis(wl.real_molecule(w), {"basis": "cc-pVDZ"})
    gd.set_basis(0, *mol["primary"].args())
    gd.set_basis(1, *mol["auxiliary"].args())
    gd.compute_two_center(False)
    mask, _ = gd.compute_schwarz_mask(1e-5)
    assert np.array_equal(mask, wl.real_mask(w))  # committed fixture of the reference's Schwarz rule
    gd.compute_B()
    return w, gd


_SD_CACHE = {}


def oracle_fock_from_device_B(gd, q_len, C, H, mask, slab):
    """F_oracle = H + sum over row slabs of the reference's screened build on B rows read back from the device."""
    key = (mask.shape[0], int(mask.sum()), hash(np.packbits(mask).tobytes()))
    sd = _SD_CACHE.get(key) or _SD_CACHE.setdefault(key, orc.ScreeningData(mask))
    P = sd.screened_indices_count
    F2 = orc.fock_build_screened_slabs(lambda q0, q1: gd.get_B_rows(0, q0, q1 - q0, P), q_len, C, sd, CORES, slab)
    return np.asfortranarray(F2 + H)


@pytest.fixture(scope="module")
def w50_default():
    """(F_gpu of the default option set, F_oracle) for bench.py's seeded C_occ and H."""
    w = wl.WORKLOADS["w50"]
    C, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
    _, gd = _w50(0)
    F = gd.fock_build(C, H)[0].copy()
    launches = gd.layout().kernel_launches_per_build
    F_orc = oracle_fock_from_device_B(gd, w.Q, C, H, wl.real_mask(w), 1200)
    gd.close()
    return w, C, H, F, F_orc, launches


def test_w50_default_kernels_match_the_oracle(w50_default):
    w, C, H, F, F_orc, launches = w50_default
    assert np.isfinite(F).all() and np.array_equal(F, F.T)
    err = float(np.abs(F - F_orc).max())
    print("real w50: max|F_gpu - F_oracle| = %.3e on max|F| = %.3f (%d launches per build)" % (err, np.abs(F).max(), launches))
    assert err < TOL


def test_w50_fixture_is_the_oracle_value(w50_default):
    """The samples bench.py asserts at every GPU count are ORACLE values (tools/make_w50_fock_fixture.py)."""
    w, C, H, F, F_orc, _ = w50_default
    fx = json.load(open(os.path.join(ROOT, "tests", "golden", "w50_fock_fixture.json")))
    idx = np.array(fx["index"])
    vals = np.array(fx["value"])
    assert len(vals) == 4096
    assert np.abs(F_orc[idx[:, 0], idx[:, 1]] - vals).max() < TOL
    assert np.abs(F[idx[:, 0], idx[:, 1]] - vals).max() < TOL
    assert abs(np.linalg.norm(F) - fx["frobenius"]) < 1e-8


def test_w50_tile_family_and_q_shards_match_the_oracle(w50_default):
    w, C, H, F, F_orc, _ = w50_default
    _, gd = _w50(1)
    F_tiles = gd.fock_build(C, H)[0].copy()
    gd.close()
    assert np.array_equal(F_tiles, F_tiles.T)
    assert np.abs(F_tiles - F_orc).max() < TOL
    # Q shards: [0, 2200) with H, [2200, 4800) without; ragged 128-row chunks on both
    parts = np.zeros_like(F)
    for r, Hr in ((range(0, 2200), H), (range(2200, 4800), None)):
        _, gd = _w50(0, shard=r)
        parts += gd.fock_build(C, Hr)[0]
        gd.close()
    assert np.abs(parts - F_orc).max() < TOL


# (workload, first auxiliary row of the slab, rows): two chunks of 128 with a ragged second one; the gly18 shard is the
# piece one of 8 GPUs owns, so its slab exercises N = 3395 (27 exchange tile rows, three orbital tiles)
SLABS = [("benzene_dimer", 700, 200), ("c20h42", 1100, 200), ("gly18_q1167", 500, 144)]


@pytest.mark.parametrize("name,q0,rows", SLABS)
@pytest.mark.parametrize("ht_mode", [0, 1])
def test_baseline_shapes_match_the_oracle_on_a_row_slab(name, q0, rows, ht_mode):
    w = wl.WORKLOADS[name]
    mask = wl.screening_mask(w)
    if ht_mode == 1 and mask is None:
        pytest.skip("dense mask: the automatic choice already is the tile family")
    C, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
    gd = jc.B200Data(w.N, w.Q, w.occ, 1)
    gd.set_shards([range(q0, q0 + rows)])
    gd.set_option("ht_mode", ht_mode)
    gd.set_mask(jc.ScreeningData(mask, w.N))
    gd.fill_synthetic_B(w.seed, w.decay_length)
    F = gd.fock_build(C, H)[0].copy()
    m = mask if mask is not None else np.ones((w.N, w.N), dtype=bool, order="F")
    # spot-check the device rows against the host twin of the generator before trusting them as the oracle's input
    keep = np.flatnonzero(np.asfortranarray(m).ravel(order="F"))  # reference pair order: column = q + N p
    P = len(keep)
    Bq = gd.get_B_rows(0, 3, 1, P)
    lib = jc.load_library()
    for k in range(0, P, max(1, P // 97)):
        c = int(keep[k])
        assert Bq[0, k] == lib.jcdf_synthetic_value(w.seed, w.Q, q0 + 3, c % w.N, c // w.N, w.decay_length)
    F_orc = oracle_fock_from_device_B(gd, rows, C, H, m, rows)
    gd.close()
    err = float(np.abs(F - F_orc).max())
    print("%s rows [%d, %d) ht_mode %d: max|F_gpu - F_oracle| = %.3e on max|F| = %.3f" % (name, q0, q0 + rows, ht_mode, err, np.abs(F).max()))
    assert np.array_equal(F, F.T)
    assert err < TOL


@pytest.mark.parametrize("name,q0,rows", [s for s in SLABS if s[0] != "benzene_dimer"])
def test_chain_shapes_with_block_screened_exchange_match_the_oracle_on_a_row_slab(name, q0, rows):
    """df_exchange_screen = true at the reference's default 10 K blocks on the chain-like benchmark shapes, where whole
    exchange tiles and 64-row halves drop out of the work lists (C20H42 shape: 21 of 66 tiles + 6 halves; gly18: 170 of 378
    + 22).  Oracle: in the computed blocks the reference's screened build on the device's own rows, in the dropped blocks the
    caller's buffer + H (calculate_K_lower_diagonal_block never writes them, ScreenedDF.jl:494-512; the oracle's own
    restatement of that equivalence is checked in tests/test_oracle_algorithms.py)."""
    w = wl.WORKLOADS[name]
    mask = wl.screening_mask(w)
    C, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
    gd = jc.B200Data(w.N, w.Q, w.occ, 1)
    gd.set_shards([range(q0, q0 + rows)])
    gd.set_option("exchange_block_screen", 10)
    gd.set_mask(jc.ScreeningData(mask, w.N))
    gd.fill_synthetic_B(w.seed, w.decay_length)
    li = gd.layout()
    nb, width, batch = orc.exchange_block_layout(w.N, 10, 10)
    todo, bsm = orc.exchange_block_screen_matrix(mask, nb, width, batch, True)
    assert (li.exchange_blocks_total, li.exchange_blocks_kept) == (len(batch), len(todo)) and len(todo) < len(batch)
    assert li.exchange_tiles_skipped > 0 and li.exchange_half_tiles_skipped > 0
    dropped = np.zeros((w.N, w.N), dtype=bool)
    dropped[:nb * width, :nb * width] = ~np.kron(bsm | bsm.T, np.ones((width, width), dtype=bool))
    F_in = np.asfortranarray(np.random.default_rng(3).standard_normal((w.N, w.N)))
    F_in = np.asfortranarray(F_in + F_in.T)
    F = F_in.copy(order="F")
    gd.fock_build(C, H, F)
    F_orc = oracle_fock_from_device_B(gd, rows, C, H, mask, rows)
    gd.close()
    F_orc[dropped] = (F_in + H)[dropped]
    err = float(np.abs(F - F_orc).max())
    print("%s rows [%d, %d) block-screened: %d of %d blocks, %d tiles + %d halves skipped of %d, max|F_gpu - F_oracle| = %.3e" % (
        name, q0, q0 + rows, len(todo), len(batch), li.exchange_tiles_skipped, li.exchange_half_tiles_skipped,
        li.exchange_tiles_total, err))
    assert np.array_equal(F, F.T)
    assert np.array_equal(F[dropped], (F_in + H)[dropped])
    assert err < TOL


@pytest.mark.slow
def test_w50_converged_energy_matches_the_oracle_scf():
    """North star: converged total energy within 1e-8 Eh.  The whole real-w50 SCF on the device with the benched (default)
    kernels against the oracle's replay of the reference's loop on the host cores, same S, H, E_nuc, mask and B
    (tools/w50_scf_parity.py; about three minutes of CPU time).  JCDF_SCF_PARITY_MOLECULES shrinks the cluster."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import w50_scf_parity
    n = int(os.environ.get("JCDF_SCF_PARITY_MOLECULES", "50"))
    out = w50_scf_parity.run(n, 50)
    print({k: v for k, v in out.items() if k != "per_iteration"})
    assert out["converged_gpu"] and out["converged_oracle"]
    assert out["iterations_gpu"] == out["iterations_oracle"]
    assert out["abs_diff"] <= 1e-8
