"""GPU parity tests proper: libjcdf (through the C-ABI, via the ctypes binding) against the CPU oracle on the same
seeded inputs.  Tolerance: 1e-10 absolute per Fock element (BASELINE.json north_star)."""
import numpy as np
import pytest

from helpers import banded_mask, jc, oracle_fock, orc, random_inputs, reference_layout_rows, wl

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _build(N, Q, occ, B, C_occ, H, mask, ndev=1, shards=None, fused=1, overlap=0, ht_mode=0, coulomb_pp=1, pp_fused=1,
           v_in_ht=1):
    gd = jc.B200Data(N, Q, occ, ndev)
    if shards is not None:
        gd.set_shards(shards)
    gd.set_option("ht_mode", ht_mode)  # 0 = chosen from the mask, 1 = tile kernel, 2 = per-function kernel
    gd.set_option("coulomb_pp", coulomb_pp)  # with the per-function kernel: Coulomb over the compact copy too
    gd.set_option("coulomb_pp_fused", pp_fused)  # ... as one cooperative single-pass kernel (0: two streaming passes)
    gd.set_option("coulomb_v_in_ht", v_in_ht)  # compact Coulomb: V from the half-transform's accumulators
    gd.set_option("coulomb_fused", fused)
    gd.set_option("coulomb_overlap", overlap)
    gd.set_mask(jc.ScreeningData(mask, N))
    shards = shards or jc.balanced_function_shards(Q, ndev)
    for d, r in enumerate(shards):
        gd.set_B(d, reference_layout_rows(B, r, mask))
    F, st = gd.fock_build(C_occ, H)
    return gd, F, st


@pytest.mark.parametrize("N,Q,occ", [(25, 96, 5), (64, 32, 8), (70, 40, 9), (33, 7, 1), (129, 50, 17), (203, 131, 27)])
@pytest.mark.parametrize("fused,ht_mode", [(1, 1), (0, 1), (2, 1), (1, 2)])
def test_dense_parity(N, Q, occ, fused, ht_mode):
    # fused = 2 is the default: with the tile family, V from the half-transform's accumulators + one J sweep
    B, C_occ, H = random_inputs(N, Q, occ, seed=N * 1000 + Q)
    F_ref = oracle_fock(B, C_occ, H, None)
    gd, F, st = _build(N, Q, occ, B, C_occ, H, None, fused=fused, ht_mode=ht_mode)
    assert np.abs(F - F_ref).max() < TOL
    assert np.array_equal(F, F.T)
    gd.close()


@pytest.mark.parametrize("fused,overlap,ht_mode,coulomb_pp",
                         [(1, 0, 1, 1), (0, 0, 1, 1), (0, 1, 1, 1), (2, 0, 1, 1), (2, 1, 1, 1), (1, 0, 2, 1), (1, 0, 2, 2), (1, 0, 2, 0),
                          (0, 0, 2, 0)])
@pytest.mark.parametrize("N,Q,occ,hw", [(70, 40, 9, 20), (203, 131, 27, 60), (160, 64, 16, 33), (300, 90, 40, 70)])
def test_screened_parity(N, Q, occ, hw, fused, overlap, ht_mode, coulomb_pp):
    """All three Coulomb schedules (single-pass cooperative kernel, two-pass sweeps, two-pass on the second stream) and
    both half-transform kernels (32 x 32 tiles with strip skipping / per-function partner lists)."""
    mask = banded_mask(N, hw, seed=N)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + 17, mask=mask, decay=hw / 2.0)
    F_ref = oracle_fock(B, C_occ, H, mask)
    # coulomb_pp: 1 = compact Coulomb as one cooperative single-pass kernel, 2 = as two streaming passes, 0 = tile space
    gd, F, st = _build(N, Q, occ, B, C_occ, H, mask, fused=fused, overlap=overlap, ht_mode=ht_mode,
                       coulomb_pp=min(coulomb_pp, 1), pp_fused=int(coulomb_pp == 1))
    base = (6 if coulomb_pp == 1 else 7) if (ht_mode == 2 and coulomb_pp) else 6 if (fused and ht_mode == 1) or fused == 1 else 7
    assert gd.layout().kernel_launches_per_build in (base, base + 1)  # + k_exchange_reduce when the contraction is sliced
    li = gd.layout()
    assert li.pairs_kept == int(mask.sum())
    assert li.pairs_tri == int(np.tril(mask).sum())
    assert li.tiles_kept <= li.tiles_total
    assert np.abs(F - F_ref).max() < TOL
    assert np.array_equal(F, F.T)
    gd.close()


def test_screened_equals_dense_algorithm_when_mask_full():
    N, Q, occ = 96, 48, 12
    mask = np.asfortranarray(np.ones((N, N), dtype=bool))
    B, C_occ, H = random_inputs(N, Q, occ, seed=5)
    F_ref = oracle_fock(B, C_occ, H, None)
    gd, F, _ = _build(N, Q, occ, B, C_occ, H, mask)
    assert np.abs(F - F_ref).max() < TOL
    gd.close()


def test_fused_coulomb_is_deterministic_and_long_rows_fall_back():
    """The cooperative kernel sums per-CTA partials in a fixed order: repeated builds are bit-identical.  A shard whose
    row segment exceeds the per-thread register segment (N = 1500 unmasked: 62 KB per CTA > 56 KB) silently uses the two-pass sweeps."""
    N, Q, occ = 203, 131, 27
    B, C_occ, H = random_inputs(N, Q, occ, seed=9)
    gd, F1, _ = _build(N, Q, occ, B, C_occ, H, None, fused=1)
    for _ in range(3):
        assert np.array_equal(gd.fock_build(C_occ, None)[0], F1)
    gd.close()
    N, Q, occ = 1500, 3, 2
    B, C_occ, H = random_inputs(N, Q, occ, seed=10)
    gd, F, _ = _build(N, Q, occ, B, C_occ, H, None, fused=1, ht_mode=1)  # tile family: its Coulomb is the one under test
    assert gd.layout().kernel_launches_per_build == 7
    assert np.abs(F - oracle_fock(B, C_occ, H, None)).max() < TOL
    gd.close()


def test_repeat_builds_and_H_none():
    """H stays resident after the first call; F_out is fully overwritten each call; builds are repeatable."""
    N, Q, occ = 70, 40, 9
    B, C_occ, H = random_inputs(N, Q, occ, seed=11)
    gd, F1, _ = _build(N, Q, occ, B, C_occ, H, None)
    F_out = np.full((N, N), np.nan, order="F")
    F2, _ = gd.fock_build(C_occ, None, F_out)
    assert F2 is F_out
    assert np.array_equal(F1, F2)
    C2 = np.asfortranarray(np.linalg.qr(np.random.default_rng(3).standard_normal((N, occ)))[0])
    F3, _ = gd.fock_build(C2, H)
    assert np.abs(F3 - oracle_fock(B, C2, H, None)).max() < TOL
    gd.close()


def test_ldc_larger_than_N():
    """C_occ passed as the leading occ columns of the full N x N coefficient matrix (DensityFitting.jl:49 slices it)."""
    N, Q, occ = 70, 40, 9
    B, C_occ, H = random_inputs(N, Q, occ, seed=12)
    big = np.asfortranarray(np.random.default_rng(1).standard_normal((N + 5, N)))
    Cfull = big[:N, :]
    Cfull[:, :occ] = C_occ
    assert Cfull.strides[1] == 8 * (N + 5)
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_mask(jc.ScreeningData(None, N))
    gd.set_B(0, reference_layout_rows(B, range(0, Q), None))
    F, _ = gd.fock_build(Cfull, H)
    assert np.abs(F - oracle_fock(B, C_occ, H, None)).max() < TOL
    gd.close()


def test_B_formation_on_device():
    """jcdf_upload_metric + jcdf_upload_three_center + jcdf_factor_and_form_B against potrf/trtri/trmm (a4)."""
    N, Q, occ = 70, 60, 9
    rng = np.random.default_rng(21)
    mask = banded_mask(N, 30, seed=2)
    T, C_occ, H = random_inputs(N, Q, occ, seed=22, mask=mask)
    A = rng.standard_normal((Q, Q))
    J = A @ A.T + Q * np.eye(Q)
    J_lower = np.asfortranarray(np.tril(J))  # only the lower triangle is valid (TwoCenterIntegrals.jl:39-41)
    sd = orc.ScreeningData(mask)
    B_ref = orc.calculate_B(J_lower, orc.dense_to_screened(T, sd))
    F_ref = orc.df_rhf_fock_build([B_ref], np.hstack([C_occ, np.zeros((N, N - occ))]), occ, H, sd, "screened")
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_mask(jc.ScreeningData(mask, N))
    gd.upload_metric(J_lower)
    gd.upload_three_center(0, reference_layout_rows(T, range(0, Q), mask))
    gd.factor_and_form_B()
    Linv = gd.get_metric_inverse()
    assert np.abs(np.tril(Linv) - orc.metric_inverse_cholesky(J_lower)).max() < 1e-12
    F, _ = gd.fock_build(C_occ, H)
    assert np.abs(F - F_ref).max() < TOL
    gd.close()


def test_not_positive_definite_metric_is_an_error():
    N, Q, occ = 33, 8, 2
    T, C_occ, H = random_inputs(N, Q, occ, seed=1)
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_mask(jc.ScreeningData(None, N))
    gd.upload_metric(np.asfortranarray(-np.eye(Q)))
    gd.upload_three_center(0, reference_layout_rows(T, range(0, Q), None))
    with pytest.raises(jc.JCDFError) as e:
        gd.factor_and_form_B()
    assert e.value.status == -5
    gd.close()


def test_call_order_errors():
    gd = jc.B200Data(33, 8, 2, 1)
    with pytest.raises(jc.JCDFError):
        gd.fock_build(np.zeros((33, 2), order="F"), None)  # no mask / B yet
    gd.set_mask(jc.ScreeningData(None, 33))
    with pytest.raises(jc.JCDFError):
        gd.fock_build(np.zeros((33, 2), order="F"), None)  # B missing
    with pytest.raises(jc.JCDFError):
        jc.B200Data(33, 8, 2, 4096)  # more devices than exist (GPUDF.jl:39-41)
    gd.close()


def test_eigh_on_device():
    N = 90
    rng = np.random.default_rng(4)
    S = rng.standard_normal((N, N))
    S = S @ S.T + N * np.eye(N)
    w, U = np.linalg.eigh(S)
    X = np.asfortranarray(U @ np.diag(w ** -0.5) @ U.T)
    F = rng.standard_normal((N, N))
    F = np.asfortranarray(0.5 * (F + F.T))
    gd = jc.B200Data(N, 8, 2, 1)
    ev, C = gd.eigh(F, X)
    ev_ref = np.linalg.eigvalsh(X.T @ F @ X)
    assert np.abs(ev - ev_ref).max() < 1e-10
    # F C = S C diag(ev)
    assert np.abs(F @ C - S @ C @ np.diag(ev)).max() < 1e-8
    gd.close()


def test_synthetic_fill_matches_host_twin():
    """jcdf_fill_synthetic_B (device) == workloads.synthetic_B_rows (host twin) through a full Fock build."""
    w = wl.WORKLOADS["small"]
    mask = wl.screening_mask(w)
    B = wl.synthetic_B_rows(w, range(0, w.Q), mask)
    C_occ, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
    F_ref = oracle_fock(B, C_occ, H, mask)
    gd = jc.B200Data(w.N, w.Q, w.occ, 1)
    gd.set_mask(jc.ScreeningData(mask, w.N))
    gd.fill_synthetic_B(w.seed, w.decay_length)
    F, _ = gd.fock_build(C_occ, H)
    assert np.abs(F - F_ref).max() < TOL
    gd.close()


def test_two_shards_on_one_device_sum_to_whole():
    """Q-sharding (a12): two handles owning disjoint Q ranges; H added by shard 0 only; partials sum to the full F."""
    N, Q, occ = 70, 41, 9
    B, C_occ, H = random_inputs(N, Q, occ, seed=31)
    F_ref = oracle_fock(B, C_occ, H, None, shards=[range(0, 20), range(20, 41)])
    parts = []
    for r, Hr in ((range(0, 20), H), (range(20, 41), None)):
        gd = jc.B200Data(N, len(r), occ, 1)
        gd.set_mask(jc.ScreeningData(None, N))
        gd.set_B(0, reference_layout_rows(B, r, None))
        parts.append(gd.fock_build(C_occ, Hr)[0].copy())
        gd.close()
    assert np.abs(parts[0] + parts[1] - F_ref).max() < TOL


def _ngpu():
    try:
        return jc.device_count()
    except Exception:
        return 0


@pytest.mark.parametrize("ndev", [2, 4, 8])
def test_multi_device_q_sharding_allreduce(ndev):
    """One process driving ndev devices: contiguous Q shards (a12), H on shard 0, ncclAllReduce of the partial Fock
    matrices inside jcdf_fock_build (replaces GPUDF.jl:267-277 + DensityFitting.jl:68-71)."""
    if _ngpu() < ndev:
        pytest.skip("needs %d GPUs" % ndev)
    N, Q, occ = 203, 131, 27
    mask = banded_mask(N, 60, seed=N)
    B, C_occ, H = random_inputs(N, Q, occ, seed=77, mask=mask, decay=30.0)
    # reference rule: contiguous auxiliary shells, remainder to the last rank (DynamicLoad.jl:160-203)
    nbas = [1] * 40 + [3] * 17 + [6] * 5 + [10] * 1
    assert sum(nbas) == Q
    shards = [jc.static_load_rank_indicies(d, ndev, nbas)[1] for d in range(ndev)]
    F_ref = oracle_fock(B, C_occ, H, mask, shards=shards)
    gd, F, st = _build(N, Q, occ, B, C_occ, H, mask, ndev=ndev, shards=shards)
    assert np.abs(F - F_ref).max() < TOL
    assert np.array_equal(F, F.T)
    F2, _ = gd.fock_build(C_occ, None)
    assert np.array_equal(F, F2)
    gd.close()


def test_multi_device_B_formation():
    """jcdf_factor_and_form_B across 2 devices: B_g = sum_{h<=g} Linv[Q_g,Q_h] T_h (GPUDF.jl:918-997)."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    N, Q, occ = 70, 60, 9
    rng = np.random.default_rng(21)
    T, C_occ, H = random_inputs(N, Q, occ, seed=23)
    A = rng.standard_normal((Q, Q))
    J_lower = np.asfortranarray(np.tril(A @ A.T + Q * np.eye(Q)))
    mask = np.asfortranarray(np.ones((N, N), dtype=bool))
    sd = orc.ScreeningData(mask)
    B_ref = orc.calculate_B(J_lower, orc.dense_to_screened(T, sd))
    F_ref = orc.df_rhf_fock_build([B_ref], np.hstack([C_occ, np.zeros((N, N - occ))]), occ, H, sd, "screened")
    gd = jc.B200Data(N, Q, occ, 2)
    shards = [range(0, 25), range(25, 60)]
    gd.set_shards(shards)
    gd.set_mask(jc.ScreeningData(None, N))
    gd.upload_metric(J_lower)
    for d, r in enumerate(shards):
        gd.upload_three_center(d, reference_layout_rows(T, r, None))
    gd.factor_and_form_B()
    F, _ = gd.fock_build(C_occ, H)
    assert np.abs(F - F_ref).max() < TOL
    gd.close()


def _scattered_mask(N, n_sites, reach, seed):
    """Functions dealt to sites in an interleaved order (like atoms in file order with tight and diffuse shells mixed):
    sparse at the pair level, but in the input order nearly every 32 x 32 tile holds a kept pair."""
    rng = np.random.default_rng(seed)
    site = rng.integers(0, n_sites, size=N)
    extent = rng.choice([0, 1, 3], size=N, p=[0.5, 0.35, 0.15])  # tight / medium / diffuse functions
    d = np.abs(site[:, None] - site[None, :])
    m = d <= np.minimum(extent[:, None], extent[None, :]) * reach + (extent[:, None] + extent[None, :] > 2)
    np.fill_diagonal(m, True)
    return np.asfortranarray(m | m.T)


@pytest.mark.parametrize("N,Q,occ,ht_mode,coulomb_pp",
                         [(180, 300, 130, 2, 1), (180, 300, 130, 2, 2), (180, 300, 130, 2, 0), (333, 257, 40, 2, 1),
                          (333, 257, 40, 0, 1), (96, 129, 81, 2, 1), (700, 520, 33, 2, 1), (700, 520, 33, 2, 2)])
def test_per_function_half_transform(N, Q, occ, ht_mode, coulomb_pp):
    """The partner-compacted path (calculate_W_screened restated per function, kernels_pp.cu) on a scattered mask that
    gets regrouped: several 128-row chunks with a ragged last one, more than one occupied-orbital tile (occ = 130),
    partner counts that are not multiples of 4, and a refresh of the compact copy when B is replaced."""
    mask = _scattered_mask(N, 19, 1, seed=N + 1)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + 5, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    gd, F, _ = _build(N, Q, occ, B, C_occ, H, mask, ht_mode=ht_mode, coulomb_pp=min(coulomb_pp, 1),
                      pp_fused=int(coulomb_pp == 1))
    assert np.abs(F - F_ref).max() < TOL
    assert np.array_equal(F, F.T)
    F2, _ = gd.fock_build(C_occ, H)  # second build reuses the compact copy
    assert np.array_equal(F2, F)
    B2, C2, H2 = random_inputs(N, Q, occ, seed=N + 6, mask=mask)
    gd.set_B(0, reference_layout_rows(B2, range(0, Q), mask))  # B replaced: the compact copy must follow
    F3, _ = gd.fock_build(C2, H2)
    assert np.abs(F3 - oracle_fock(B2, C2, H2, mask)).max() < TOL
    gd.close()


@pytest.mark.parametrize("N,Q,occ,overlap", [(180, 300, 130, 0), (333, 257, 40, 0), (333, 257, 40, 1), (96, 129, 81, 0),
                                             (700, 520, 33, 0), (64, 5, 3, 0)])
def test_coulomb_v_folded_into_half_transform(N, Q, occ, overlap):
    """Compact Coulomb with V[Q] = sum_p sum_i C[p,i] W[Q,i,p] taken from the per-function half-transform's accumulators
    (default) against the density + separate V pass: both within tolerance of the oracle, two kernel launches fewer,
    repeat builds bit-identical (per-item partials are reduced in a fixed order), several orbital tiles (occ = 130)."""
    mask = _scattered_mask(N, 19, 1, seed=N + 2)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + 8, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    out = []
    for v_in_ht in (1, 0):
        gd, F, _ = _build(N, Q, occ, B, C_occ, H, mask, ht_mode=2, coulomb_pp=1, pp_fused=0, overlap=overlap, v_in_ht=v_in_ht)
        assert gd.layout().kernel_launches_per_build in ((7, 8) if v_in_ht else (9, 10))
        assert np.abs(F - F_ref).max() < TOL
        assert np.array_equal(F, F.T)
        assert np.array_equal(gd.fock_build(C_occ, None)[0], F)
        out.append(F)
        gd.close()
    assert np.abs(out[0] - out[1]).max() < 1e-12


@pytest.mark.parametrize("occ", [1, 7, 8, 9, 24, 25, 41, 81, 96, 97, 103, 104, 130, 193])
@pytest.mark.parametrize("pp_layout", [0, 1, 2])
def test_per_function_warp_layouts(occ, pp_layout):
    """All consumer-warp arrangements of the per-function kernel (2 x 4 warps with the orbitals of an item padded to 32,
    8 x 1 warps with 8-wide tiles only, two CTAs per SM of 2 x 2 warps with items padded to 16) over occupied counts that hit every ragged case: fewer than one tile, exact
    multiples of 8 / 32 / 96, one orbital over (97 = 13 tiles -> two items of 7 with 6 valid in the second), several
    items (193), and a Q range with a ragged last chunk of 9 rows (one warp busy in the 8 x 1 arrangement)."""
    N, Q = max(100, occ + 10), 137
    mask = _scattered_mask(N, 11, 1, seed=occ)
    B, C_occ, H = random_inputs(N, Q, occ, seed=occ + 31, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_option("ht_mode", 2)
    gd.set_option("pp_layout", pp_layout)
    gd.set_mask(jc.ScreeningData(mask, N))
    gd.set_B(0, reference_layout_rows(B, range(0, Q), mask))
    F, _ = gd.fock_build(C_occ, H)
    assert np.abs(F - F_ref).max() < TOL
    assert np.array_equal(F, F.T)
    gd.close()


@pytest.mark.parametrize("N,Q,occ", [(128, 40, 12), (256, 130, 30), (40, 5, 3), (130, 128, 64), (130, 188, 12), (90, 125, 30)])
@pytest.mark.parametrize("ht_mode", [1, 2])
def test_tile_edge_sizes(N, Q, occ, ht_mode):
    """Sizes on the edges of the blockings: N an exact multiple of the 128-wide exchange tile (no clipping), one tile only
    (diagonal rectangles alone), Q below / exactly one 128-row chunk, occ = 64 (two full orbital tiles of 32), a last
    chunk of 60 / 125 rows (the last 8-row tile of a warp partly beyond Q: stores must stay guarded)."""
    mask = _scattered_mask(N, 7, 1, seed=N + Q)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + Q + 1, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    gd, F, _ = _build(N, Q, occ, B, C_occ, H, mask, ht_mode=ht_mode)
    assert np.abs(F - F_ref).max() < TOL
    assert np.array_equal(F, F.T)
    gd.close()


def test_isolated_functions_and_empty_shard():
    """Functions whose only kept partner is themselves (partner list of 1, padded to 4 rows) and a device-less corner:
    a shard of zero auxiliary rows next to a full one on the same handle layout (q_len = 0 must launch nothing)."""
    N, Q, occ = 97, 33, 11
    mask = _scattered_mask(N, 11, 1, seed=2)
    iso = [3, 50, 96]
    mask[iso, :] = False
    mask[:, iso] = False
    mask[iso, iso] = True
    mask = np.asfortranarray(mask)
    B, C_occ, H = random_inputs(N, Q, occ, seed=77, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    for mode in (1, 2):
        gd, F, _ = _build(N, Q, occ, B, C_occ, H, mask, ht_mode=mode)
        assert np.abs(F - F_ref).max() < TOL
        gd.close()
    # empty shard: H only
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_shards([range(0, 0)])
    gd.set_option("ht_mode", 2)
    gd.set_mask(jc.ScreeningData(mask, N))
    gd.set_B(0, np.zeros((0, int(mask.sum())), order="F"))
    F, _ = gd.fock_build(C_occ, H)
    assert np.abs(F - H).max() == 0.0
    gd.close()


def test_compact_copy_that_does_not_fit_falls_back_to_tiles():
    """The partner-compacted second copy of B is optional: when it exceeds the cap (or cudaMalloc refuses) the automatic
    choice silently uses the tile family; forcing the per-function family is then an error, never a wrong result."""
    N, Q, occ = 200, 150, 21
    mask = _scattered_mask(N, 23, 1, seed=N)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + 3, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_option("pp_max_bytes", 1024)
    gd.set_mask(jc.ScreeningData(mask, N))
    gd.set_B(0, reference_layout_rows(B, range(0, Q), mask))
    F, _ = gd.fock_build(C_occ, H)
    assert np.abs(F - F_ref).max() < TOL
    assert gd.layout().kernel_launches_per_build in (6, 7, 8)  # tile family (fused or two-pass Coulomb, +- slice reduce)
    gd.close()
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_option("pp_max_bytes", 1024)
    gd.set_option("ht_mode", 2)
    with pytest.raises(jc.JCDFError):
        gd.set_mask(jc.ScreeningData(mask, N))
    gd.close()


def test_half_transform_kernels_agree_on_sharded_run():
    """Two shards on one device, per-function kernel on each, against the tile kernel: same Fock matrix to 1e-10."""
    N, Q, occ = 150, 200, 33
    mask = _scattered_mask(N, 13, 1, seed=9)
    B, C_occ, H = random_inputs(N, Q, occ, seed=10, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    out = {}
    for mode in (1, 2):
        acc = np.zeros((N, N))
        for r, Hr in ((range(0, 135), H), (range(135, 200), None)):
            gd = jc.B200Data(N, len(r), occ, 1)
            gd.set_option("ht_mode", mode)
            gd.set_mask(jc.ScreeningData(mask, N))
            gd.set_B(0, reference_layout_rows(B, r, mask))
            acc += gd.fock_build(C_occ, Hr)[0]
            gd.close()
        out[mode] = acc
        assert np.abs(acc - F_ref).max() < TOL
    assert np.abs(out[1] - out[2]).max() < TOL


@pytest.mark.parametrize("regroup", [1, 0])
@pytest.mark.parametrize("N,Q,occ", [(200, 48, 21), (333, 37, 40)])
def test_regrouped_layout_is_exact(N, Q, occ, regroup):
    """Mask-driven regrouping of the basis functions (internal order only): same Fock matrix to 1e-10, fewer stored
    tiles; uploads, the packed round trip (jcdf_get_B) and the epilogue all go through the permutation."""
    mask = _scattered_mask(N, 23, 1, seed=N)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + 3, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_option("regroup", regroup)
    gd.set_mask(jc.ScreeningData(mask, N))
    rows = reference_layout_rows(B, range(0, Q), mask)
    gd.set_B(0, rows)
    li = gd.layout()
    assert np.array_equal(gd.get_B(0, Q, li.pairs_kept), rows)  # bit-exact round trip through the packed layout
    F, _ = gd.fock_build(C_occ, H)
    assert np.abs(F - F_ref).max() < TOL
    assert np.array_equal(F, F.T)
    tiles = li.tiles_kept
    gd.close()
    if regroup:
        gd0 = jc.B200Data(N, Q, occ, 1)
        gd0.set_option("regroup", 0)
        gd0.set_mask(jc.ScreeningData(mask, N))
        assert tiles < gd0.layout().tiles_kept
        gd0.close()


def test_regrouped_synthetic_fill_matches_host_twin():
    """jcdf_fill_synthetic_B writes through the permutation: the packed shard read back in reference order equals the
    host twin value for value."""
    w = wl.WORKLOADS["small"]
    mask = _scattered_mask(w.N, 17, 1, seed=4)
    gd = jc.B200Data(w.N, w.Q, w.occ, 1)
    gd.set_mask(jc.ScreeningData(mask, w.N))
    gd.fill_synthetic_B(w.seed, w.decay_length)
    P = int(mask.sum())
    got = gd.get_B(0, w.Q, P)
    ref = reference_layout_rows(wl.synthetic_B_rows(w, range(0, w.Q), mask), range(0, w.Q), mask)
    assert np.array_equal(got, ref)
    gd.close()


@pytest.mark.parametrize("ht_mode", [1, 2])
def test_graph_replay_is_bit_identical_to_eager_launches(ht_mode):
    """The per-build kernel sequence captured into a CUDA graph (default) against eager launches: same kernels, same order,
    bit-identical Fock matrices; the replay counter shows the graph really ran; stage times stay available (the timing
    events are external event-record nodes of the graph)."""
    N, Q, occ = 203, 131, 27
    mask = _scattered_mask(N, 19, 1, seed=77)
    B, C_occ, H = random_inputs(N, Q, occ, seed=78, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    res = {}
    for graph in (1, 0):
        gd = jc.B200Data(N, Q, occ, 1)
        gd.set_option("ht_mode", ht_mode)
        gd.set_option("graph", graph)
        gd.set_mask(jc.ScreeningData(mask, N))
        gd.set_B(0, reference_layout_rows(B, range(0, Q), mask))
        Fs = [gd.fock_build(C_occ, H)[0].copy() for _ in range(3)]
        C2 = np.asfortranarray(C_occ[:, ::-1])  # another density through the same graph
        F2, st = gd.fock_build(C2, None)
        assert np.array_equal(Fs[0], Fs[1]) and np.array_equal(Fs[1], Fs[2])
        assert np.abs(Fs[0] - F_ref).max() < TOL
        assert gd.layout().graph_replays == (4 if graph else 0)
        assert st["W"] > 0 and st["K"] > 0 and st["total"] > 0
        res[graph] = (Fs[0], F2.copy())
        gd.close()
    assert np.array_equal(res[1][0], res[0][0]) and np.array_equal(res[1][1], res[0][1])


def test_refused_options_leave_the_handle_unchanged():
    N, Q, occ = 70, 40, 9
    gd = jc.B200Data(N, Q, occ, 1)
    for name, value in (("ht_mode", 3), ("coulomb_fused", -1), ("pp_layout", 3), ("nonsense", 1), ("graph", 2)):
        with pytest.raises(jc.JCDFError):
            gd.set_option(name, value)
    mask = banded_mask(N, 20, seed=N)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + 1, mask=mask)
    gd.set_mask(jc.ScreeningData(mask, N))
    for name in ("ht_mode", "regroup", "exchange_slices", "exchange_balance", "pp_layout", "pp_order", "pp_max_bytes"):
        with pytest.raises(jc.JCDFError):
            gd.set_option(name, 1)  # too late: the layout exists
    gd.set_B(0, reference_layout_rows(B, range(0, Q), mask))
    assert np.abs(gd.fock_build(C_occ, H)[0] - oracle_fock(B, C_occ, H, mask)).max() < TOL
    with pytest.raises(jc.JCDFError):
        gd.set_shards([range(0, 10)])  # after jcdf_set_mask
    gd.close()
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_option("exchange_slices", 3)  # a forced split of the exchange contraction index
    gd.set_option("pp_order", 1)
    gd.set_mask(jc.ScreeningData(mask, N))
    gd.set_B(0, reference_layout_rows(B, range(0, Q), mask))
    assert np.abs(gd.fock_build(C_occ, H)[0] - oracle_fock(B, C_occ, H, mask)).max() < TOL
    gd.close()


@pytest.mark.parametrize("ht_mode", [2, 1])
@pytest.mark.parametrize("N,Q,occ,pp_layout", [(180, 300, 130, 0), (96, 129, 81, 1), (333, 257, 40, 0)])
def test_register_split_variant_is_bit_identical(N, Q, occ, pp_layout, ht_mode):
    """Both half-transform kernels with the setmaxnreg register split (12 warps, default) against the 9-warp kernels at the
    launch cap: same arithmetic in the same order, so W, V and F agree bit for bit; both match the oracle."""
    mask = _scattered_mask(N, 19, 1, seed=N + 4)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + 9, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    out = []
    for bigreg in (1, 0):
        gd = jc.B200Data(N, Q, occ, 1)
        gd.set_option("ht_mode", ht_mode)
        gd.set_option("pp_layout", pp_layout)
        gd.set_option("pp_bigreg", bigreg)
        gd.set_option("exchange_producer", bigreg)  # the exchange kernel's dedicated-producer variant likewise
        gd.set_mask(jc.ScreeningData(mask, N))
        gd.set_B(0, reference_layout_rows(B, range(0, Q), mask))
        F = gd.fock_build(C_occ, H)[0].copy()
        assert np.abs(F - F_ref).max() < TOL
        F_again = gd.fock_build(C_occ, H)[0]
        out.append(F)
        gd.close()
        if ht_mode == 2:
            assert np.array_equal(F_again, F)
        else:  # the tile family's two-pass Coulomb adds row slabs with atomics on short rows: equal to rounding only
            assert np.abs(F_again - F).max() < 1e-12
    if ht_mode == 2:
        assert np.array_equal(out[0], out[1])
    else:
        assert np.abs(out[0] - out[1]).max() < 1e-12


def test_registered_output_buffer():
    """register_output = 1 (what julia/B200DF.jl sets): a pageable F_out that arrives twice is page-locked in place; results
    are unchanged, a different buffer afterwards works, and destroy releases the registration while the array is alive."""
    N, Q, occ = 160, 64, 16
    mask = banded_mask(N, 33, seed=N)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + 1, mask=mask)
    F_ref = oracle_fock(B, C_occ, H, mask)
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_option("register_output", 1)
    gd.set_mask(jc.ScreeningData(mask, N))
    gd.set_B(0, reference_layout_rows(B, range(0, Q), mask))
    F = np.empty((N, N), order="F")
    for _ in range(4):  # the second call registers the buffer
        F[:] = 0.0
        gd.fock_build(C_occ, H, F)
        assert np.abs(F - F_ref).max() < TOL
    G = np.empty((N, N), order="F")
    gd.fock_build(C_occ, H, G)
    gd.fock_build(C_occ, H, G)
    assert np.array_equal(F, G)
    gd.close()
    F[:] = 1.0  # still ordinary memory
    assert F.sum() == N * N
