"""The oracle's two restated algorithms against an independent statement of the maths
F = H + 2 J[D] - K[D] (einsum), edge cases of the screened bookkeeping, B formation and iteration()."""
import numpy as np
import pytest

from helpers import banded_mask, oracle_fock, orc, random_inputs


def _einsum_fock(B, C_occ, H):
    D = C_occ @ C_occ.T
    V = np.einsum("Qrs,rs->Q", B, D)
    J = np.einsum("Qpq,Q->pq", B, V)
    W = np.einsum("Qpr,ri->Qip", B, C_occ)
    K = np.einsum("Qip,Qiq->pq", W, W)
    return H + 2.0 * J - K


@pytest.mark.parametrize("N,Q,occ", [(25, 96, 5), (33, 7, 1), (101, 20, 9), (127, 31, 13)])
def test_dense_algorithm_is_the_maths(N, Q, occ):
    B, C_occ, H = random_inputs(N, Q, occ, seed=N)
    assert np.abs(oracle_fock(B, C_occ, H, None) - _einsum_fock(B, C_occ, H)).max() < 1e-11


@pytest.mark.parametrize("N,Q,occ,hw", [(70, 40, 9, 20), (101, 17, 6, 40), (130, 25, 11, 129), (203, 31, 27, 60)])
def test_screened_algorithm_is_the_maths_on_masked_B(N, Q, occ, hw):
    """N >= 100 exercises the 10-block K tiling incl. the N % 10 remainder strip (ScreenedDF.jl:619-640)."""
    mask = banded_mask(N, hw, seed=N)
    B, C_occ, H = random_inputs(N, Q, occ, seed=N + 1, mask=mask)
    F = oracle_fock(B, C_occ, H, mask)
    assert np.abs(F - _einsum_fock(B, C_occ, H)).max() < 1e-11
    assert np.array_equal(F, F.T)
    Ft = orc.fock_build_screened_rank_threaded(orc.dense_to_screened(B, orc.ScreeningData(mask)), C_occ,
                                               orc.ScreeningData(mask), 4) + H
    assert np.abs(Ft - F).max() < 1e-11


def test_rank_sum_equals_whole_and_H_added_once():
    N, Q, occ = 70, 41, 9
    B, C_occ, H = random_inputs(N, Q, occ, seed=3)
    whole = oracle_fock(B, C_occ, H, None)
    parts = oracle_fock(B, C_occ, H, None, shards=[range(0, 13), range(13, 13), range(13, 41)])  # one empty rank
    assert np.abs(whole - parts).max() < 1e-11


def test_screening_metadata_ranges():
    mask = np.array([[1, 1, 0, 1], [1, 1, 1, 0], [0, 1, 1, 1], [1, 0, 1, 1]], dtype=bool)
    sd = orc.ScreeningData(mask)
    assert sd.screened_indices_count == 12
    assert list(sd.sparse_p_start_indices) == [1, 4, 7, 10]
    assert list(sd.non_screened_p_indices_count) == [3, 3, 3, 3]
    assert sd.non_zero_ranges[0] == [(1, 2), (4, 4)] and sd.non_zero_sparse_ranges[0] == [(1, 2), (3, 3)]
    with pytest.raises(ValueError):
        orc.ScreeningData(np.array([[1, 0], [0, 0]], dtype=bool))


def test_exchange_block_layout_rules():
    assert orc.exchange_block_layout(99)[:2] == (1, 99)          # N < 100 -> single block (ScreenedDF.jl:396-400)
    nb, w, batch = orc.exchange_block_layout(1250)
    assert (nb, w, len(batch)) == (10, 125, 55)
    assert orc.exchange_block_layout(203, 4)[:2] == (4, 50)


def test_B_formation_is_cholesky_inverse_not_symmetric_root():
    rng = np.random.default_rng(0)
    Q, P = 30, 50
    A = rng.standard_normal((Q, Q))
    J = A @ A.T + Q * np.eye(Q)
    T = np.asfortranarray(rng.standard_normal((Q, P)))
    B = orc.calculate_B(np.tril(J), T)       # only the lower triangle may be read
    L = np.linalg.cholesky(J)
    assert np.abs(B - np.linalg.solve(L, T)).max() < 1e-12
    assert np.abs(B.T @ B - T.T @ np.linalg.solve(J, T)).max() < 1e-11   # the fitted 4-index tensor is the invariant
    rows = orc.calculate_B_rows(orc.metric_inverse_cholesky(np.tril(J)), T, range(7, 19))
    assert np.abs(rows - B[7:19]).max() < 1e-12
    with pytest.raises(np.linalg.LinAlgError):
        orc.calculate_B(-np.eye(4), np.zeros((4, 3), order="F"))


def test_iteration_energy_and_density():
    rng = np.random.default_rng(1)
    N, nocc = 12, 3
    A = rng.standard_normal((N, N)); S = A @ A.T + N * np.eye(N)
    w, U = np.linalg.eigh(S); X = U @ np.diag(w ** -0.5) @ U.T
    F = rng.standard_normal((N, N)); F = F + F.T
    H = rng.standard_normal((N, N)); H = H + H.T
    E, ev, C, D = orc.iteration(F, H, X, nocc)
    assert np.abs(C.T @ S @ C - np.eye(N)).max() < 1e-10
    assert np.abs(F @ C - S @ C @ np.diag(ev)).max() < 1e-9
    assert abs(np.trace(D @ S) - 2 * nocc) < 1e-10
    assert abs(E - 0.5 * np.sum(D * (F + H))) < 1e-12


def test_slab_sum_of_the_threaded_build_is_the_whole_build():
    """fock_build_screened_slabs (used at full size, where B is read back from the device slab by slab) against the
    one-shot threaded build and the einsum statement; ragged last slab."""
    N, Q, occ = 130, 37, 11
    mask = banded_mask(N, 50, seed=5)
    B, C_occ, H = random_inputs(N, Q, occ, seed=8, mask=mask)
    sd = orc.ScreeningData(mask)
    Bs = orc.dense_to_screened(B, sd)
    tm = {}
    F = orc.fock_build_screened_slabs(lambda q0, q1: Bs[q0:q1], Q, C_occ, sd, 3, 10, tm) + H
    assert np.abs(F - _einsum_fock(B, C_occ, H)).max() < 1e-11
    assert np.abs(F - (orc.fock_build_screened_rank_threaded(Bs, C_occ, sd, 3) + H)).max() < 1e-11
    assert set(tm) == {"W", "K", "J"}


def test_block_screened_exchange_of_the_reference_depends_on_the_iteration_history():
    """Row a7 with df_exchange_screen = true (ScreenedDF.jl:433-449, 459-545), restated: the blocks whose mask sub-block is
    empty are neither computed NOR written, so two_electron_fock keeps its previous content there and the new J is added on
    top.  The same (B, C) therefore gives different Fock matrices depending on what the buffer held — libjcdf reproduces it
    with F_out as an in/out argument (option "exchange_block_screen", tests/test_gpu_block_screened_exchange.py)."""
    N, Q, occ = 120, 16, 7
    mask = banded_mask(N, 18, seed=3, speckle=False)
    B, C_occ, H = random_inputs(N, Q, occ, seed=4, mask=mask)
    sd = orc.ScreeningData(mask)
    Bs = orc.dense_to_screened(B, sd)
    n_blocks, width, batch = orc.exchange_block_layout(N, 10, 10)
    todo, block_screen = orc.exchange_block_screen_matrix(mask, n_blocks, width, batch, True)
    assert 0 < len(todo) < len(batch) and not block_screen[n_blocks - 1, 0]
    todo_all, _ = orc.exchange_block_screen_matrix(mask, n_blocks, width, batch, False)
    assert todo_all == list(range(len(batch)))
    C_T = np.asfortranarray(C_occ.T)
    W = orc.calculate_W_screened(Bs, C_T, sd)

    def build(fock):
        orc.calculate_K_lower_diagonal_block(W, fock, n_blocks, width, batch, todo)
        orc.calculate_coulomb_screened(Bs, C_T, sd, fock)
        return fock

    first = build(np.zeros((N, N), order="F"))           # iteration 1: the buffer starts as zeros (DensityFitting.jl:45)
    stale = np.asfortranarray(first + H)                  # what the buffer holds when iteration 2 starts (F .+= H, :62-65)
    second = build(stale.copy())                          # same B, same C
    dropped = ~np.kron(block_screen | block_screen.T, np.ones((width, width), dtype=bool))
    assert np.abs(second - first)[~dropped].max() < 1e-12          # computed blocks: overwritten, identical
    assert np.abs(second - first)[dropped].max() > 1e-3            # dropped blocks: previous values leak through
    # with every block computed the build is a function of (B, C) again and equals the default algorithm
    full = np.zeros((N, N), order="F")
    orc.calculate_K_lower_diagonal_block(W, full, n_blocks, width, batch, todo_all)
    orc.calculate_coulomb_screened(Bs, C_T, sd, full)
    assert np.abs(full - orc.fock_build_screened_rank(Bs, C_occ, sd, 10)).max() < 1e-12


def test_block_screened_build_over_ranks_keeps_rank_count_copies_of_the_old_matrix():
    """df_rhf_fock_build! with df_exchange_screen over n emulated ranks (DensityFitting.jl:62-71 around ScreenedDF.jl:459-545):
    in the computed blocks the result is the default algorithm's; a dropped block comes back as n * old + H, because every
    rank keeps its own copy of the old matrix there, rank 0 adds H and the all-reduce sums the ranks."""
    N, Q, occ = 130, 18, 6
    mask = banded_mask(N, 20, seed=5)
    B, C_occ, H = random_inputs(N, Q, occ, seed=6, mask=mask)
    sd = orc.ScreeningData(mask)
    Cfull = np.zeros((N, N), order="F")
    Cfull[:, :occ] = C_occ
    n_blocks, width, batch = orc.exchange_block_layout(N, 0, 10)
    assert (n_blocks, width) == (10, 13)
    _, block_screen = orc.exchange_block_screen_matrix(mask, n_blocks, width, batch, True)
    dropped = np.zeros((N, N), dtype=bool)
    dropped[:n_blocks * width, :n_blocks * width] = ~np.kron(block_screen | block_screen.T, np.ones((width, width), dtype=bool))
    assert dropped.any() and not dropped[:, n_blocks * width:].any()
    old = np.asfortranarray(np.random.default_rng(9).standard_normal((N, N)))
    old = np.asfortranarray(old + old.T)
    for shards in ([range(0, Q)], [range(0, 7), range(7, 7), range(7, Q)]):
        Bs = [orc.dense_to_screened(np.asfortranarray(B[r.start:r.stop]), sd) for r in shards]
        F_default = orc.df_rhf_fock_build(Bs, Cfull, occ, H, sd, "screened")
        F = orc.df_rhf_fock_build(Bs, Cfull, occ, H, sd, "screened", df_screen_exchange=True, two_electron_fock=old)
        assert np.abs(F - F_default)[~dropped].max() < 1e-12
        assert np.abs(F - (len(shards) * old + H))[dropped].max() < 1e-12
        F1 = orc.df_rhf_fock_build(Bs, Cfull, occ, H, sd, "screened", df_screen_exchange=True)  # iteration 1: zeros
        assert np.abs(F1 - H)[dropped].max() == 0.0
