"""CPU oracle for the DF-RHF Fock build (TEST INFRASTRUCTURE — never shipped, never on the product path).

A numpy/OpenBLAS restatement of the reference's two CPU algorithms for the density-fitted RHF Fock build,
operating on a given (B or T+metric, mask, C_occ, H).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module.

All file:line citations are relative to /root/reference (j-hayes/JuliaChem.jl @ 2064729):

  dense algorithm       src/rhf/energy/DensityFitting/DensityFitting.jl:185-224   (calculate_coulomb!, calculate_exchange!)
  B formation           src/rhf/energy/DensityFitting/DensityFitting.jl:128-183   (calculate_B!: potrf, trtri, trmm)
  screened algorithm    src/rhf/energy/DensityFitting/ScreenedDF.jl:227-378, 548-641
  screening metadata    src/rhf/energy/DensityFitting/ScreenedDF.jl:16-77, 385-457
  pair compaction       src/rhf/energy/DensityFitting/SchwarzScreening.jl:72-81
  Q sharding            src/rhf/energy/DensityFitting/DynamicLoad.jl:160-203
  dispatcher            src/rhf/energy/DensityFitting/DensityFitting.jl:23-76      (F += H on rank 0, Allreduce)
  SCF iteration         src/rhf/energy/SCF.jl:1072-1125                            (iteration())

Parity pin: the reference has no element-wise J/K/F test.  This restatement is pinned against the
reference's own logged run water_ccpvdz_out.log (per-iteration DF-RHF energies and the converged energy
-75.9911548795 Eh) through oracle/scf_replay.py + oracle/gaussian_integrals.py; see tests/test_oracle_water_golden.py.

Conventions: arrays mirror Julia's column-major storage (numpy order='F'); index maps keep the reference's
1-based values with 0 meaning "screened"; python loop variables are 0-based.
"""
from __future__ import annotations

import numpy as np
import scipy.linalg as sla
from scipy.linalg import blas as sblas
from scipy.linalg import lapack as slapack


# --------------------------------------------------------------------------------------------------------------
# a12  Q sharding (DynamicLoad.jl:160-203)
# --------------------------------------------------------------------------------------------------------------
def get_df_static_shell_indices(n_aux_shells: int, n_ranks: int, rank: int) -> range:
    """0-based half-open shell range of `rank` (DynamicLoad.jl:160-171): n_shells // n_ranks contiguous shells
    per rank, the last rank also takes the remainder."""
    n_indices = n_aux_shells // n_ranks
    begin = n_indices * rank
    end = begin + n_indices
    if rank == n_ranks - 1:
        end = n_aux_shells
    return range(begin, end)


def static_load_rank_indicies(rank: int, n_ranks: int, aux_shell_nbas):
    """(shell range, function range) of `rank`, 0-based half-open (DynamicLoad.jl:174-203).
    aux_shell_nbas[s] = number of (Cartesian) functions of auxiliary shell s; functions are laid out shell after
    shell (Shell.pos), so a contiguous shell range is a contiguous function range."""
    aux_shell_nbas = np.asarray(aux_shell_nbas, dtype=np.int64)
    pos = np.concatenate([[0], np.cumsum(aux_shell_nbas)])
    shells = get_df_static_shell_indices(len(aux_shell_nbas), n_ranks, rank)
    if len(shells) == 0:
        return shells, range(0, 0)
    return shells, range(int(pos[shells.start]), int(pos[shells.stop]))


# --------------------------------------------------------------------------------------------------------------
# a5  pair compaction (SchwarzScreening.jl:72-81) and a6 screening metadata (ScreenedDF.jl:16-77)
# --------------------------------------------------------------------------------------------------------------
def schwarz_threshold(sigma: float, max_P_P: float) -> float:
    """tau = sigma^2 / max_P (P|P)  (SchwarzScreening.jl:22; get_max_P_P :86-95)."""
    return sigma * sigma / max_P_P


def schwarz_function_mask(diag_eri_by_shell_pair, shell_pos, shell_nbas, threshold: float):
    """Shell-pair + function-pair Schwarz mask (SchwarzScreening.jl:23-71).

    diag_eri_by_shell_pair[(ish, jsh)] (ish >= jsh) is the full (nmu*nnu) x (nmu*nnu) block (mu nu|mu' nu') of the
    shell quartet (ish jsh|ish jsh) flattened as the reference indexes it (:62): element
    [nu + nnu*mu + nnu*nmu*nu' + nnu*nmu*nnu*mu'].  A shell pair is kept iff |sum of the whole block| >= threshold
    (:46-50, a plain sum, not a max); inside kept shell pairs a function pair is kept iff |(mu nu|mu nu)| >=
    threshold (:62-65)."""
    nsh = len(shell_pos)
    N = int(shell_pos[-1] + shell_nbas[-1])
    shell_mask = np.zeros((nsh, nsh), dtype=bool)
    fn_mask = np.zeros((N, N), dtype=bool, order="F")
    for ish in range(nsh):
        for jsh in range(ish + 1):
            blk = np.asarray(diag_eri_by_shell_pair[(ish, jsh)], dtype=np.float64).ravel()
            nmu, nnu = int(shell_nbas[ish]), int(shell_nbas[jsh])
            keep = not (abs(blk.sum()) < threshold)
            shell_mask[ish, jsh] = shell_mask[jsh, ish] = keep
            if not keep:
                continue
            for m in range(nmu):
                for n in range(nnu):
                    idx = n + nnu * m + nnu * nmu * n + nnu * nmu * nnu * m
                    k = not (abs(blk[idx]) < threshold)
                    fn_mask[shell_pos[ish] + m, shell_pos[jsh] + n] = k
                    fn_mask[shell_pos[jsh] + n, shell_pos[ish] + m] = k
    return shell_mask, fn_mask


def sparse_pq_index_map(fn_mask: np.ndarray) -> np.ndarray:
    """sparse_pq_index_map[q, p] = running 1-based index over kept pairs, OUTER loop p, INNER loop q
    (SchwarzScreening.jl:72-81); 0 = screened.  Returned as int64 (N, N) column-major like the Julia array."""
    N = fn_mask.shape[0]
    m = np.zeros((N, N), dtype=np.int64, order="F")
    flat_keep = np.asfortranarray(fn_mask).ravel(order="F")  # order: q fastest within p == loop order
    m_flat = m.reshape(-1, order="F")
    m_flat[flat_keep] = np.arange(1, int(flat_keep.sum()) + 1, dtype=np.int64)
    return m


class ScreeningData:
    """Subset of the reference's ScreeningData (src/shared/SCFData.jl:1-17) the hot path reads."""

    def __init__(self, fn_mask: np.ndarray):
        fn_mask = np.asfortranarray(fn_mask.astype(bool))
        N = fn_mask.shape[0]
        self.basis_function_screen_matrix = fn_mask
        self.sparse_pq_index_map = sparse_pq_index_map(fn_mask)
        self.screened_indices_count = int(fn_mask.sum())  # ScreenedDF.jl:29
        # ScreenedDF.jl:34-44: first kept row of column p -> its sparse index; kept count per column
        self.sparse_p_start_indices = np.zeros(N, dtype=np.int64)
        self.non_screened_p_indices_count = fn_mask.sum(axis=0).astype(np.int64)
        for p in range(N):
            rows = np.flatnonzero(fn_mask[:, p])
            if len(rows) == 0:
                raise ValueError("column %d of the screen matrix is empty (the reference loops forever here, "
                                 "ScreenedDF.jl:36-42)" % p)
            self.sparse_p_start_indices[p] = self.sparse_pq_index_map[rows[0], p]
        # ScreenedDF.jl:48-70 contiguous kept row ranges (1-based inclusive, like the reference's UnitRange)
        self.non_zero_ranges = []
        self.non_zero_sparse_ranges = []
        for p in range(N):
            col = fn_mask[:, p]
            rr, sr = [], []
            start = end = 0
            nz = 1
            for r in range(1, N + 1):
                if col[r - 1]:
                    if start == 0:
                        start = r
                    end = r
                if start != 0 and ((not col[r - 1]) or r == N):
                    rr.append((start, end))
                    ln = end - start + 1
                    sr.append((nz, nz + ln - 1))
                    nz += ln
                    start = end = 0
            self.non_zero_ranges.append(rr)
            self.non_zero_sparse_ranges.append(sr)


def exchange_block_layout(N: int, df_exchange_n_blocks: int = 0, default_n_blocks: int = 10):
    """a7 (ScreenedDF.jl:385-417): returns (n_blocks, K_block_width, lower-triangular block list (1-based))."""
    n_blocks = df_exchange_n_blocks if df_exchange_n_blocks != 0 else default_n_blocks
    if N < 100:
        width, n_blocks = N, 1
    else:
        width = N // n_blocks
    batch = [(i, j) for i in range(1, n_blocks + 1) for j in range(1, i + 1)]
    return n_blocks, width, batch


def exchange_block_screen_matrix(fn_mask: np.ndarray, n_blocks: int, width: int, batch, df_screen_exchange: bool):
    """Second half of a7 (ScreenedDF.jl:425-456): blocks_to_calculate (0-based positions in `batch`) and the Bool
    block_screen_matrix.  With df_exchange_screen a lower-triangular block (pp, qq) is kept iff its width x width
    sub-block of basis_function_screen_matrix holds any kept pair (:433-441); without it every block is kept (:443-446)."""
    block_screen = np.zeros((n_blocks, n_blocks), dtype=bool)
    todo = []
    for idx, (pp, qq) in enumerate(batch):
        keep = True
        if df_screen_exchange:
            keep = bool(fn_mask[(pp - 1) * width:pp * width, (qq - 1) * width:qq * width].sum() != 0)
        if keep:
            todo.append(idx)
            block_screen[pp - 1, qq - 1] = True
    return todo, block_screen


def calculate_K_lower_diagonal_block(W: np.ndarray, fock: np.ndarray, n_blocks: int, width: int, batch, blocks_to_calculate):
    """ScreenedDF.jl:459-545 (df_exchange_screen = true): like the no-screen variant, but ONLY the blocks in
    blocks_to_calculate are computed and written (:494-510).  The GEMM's beta = 0 write is the only thing that ever
    overwrites scf_data.two_electron_fock, so a dropped block keeps whatever the matrix held — after the first iteration
    that is the previous iteration's F (H, J, K and all, after the caller's DIIS and damping: SCF.jl:471-508 mutate the very
    array df_rhf_fock_build! returned).  calculate_coulomb_screened adds nothing there either: every pair of a dropped block
    is screened (:367-378).  libjcdf's option "exchange_block_screen" reproduces this with F_out as an in/out argument."""
    Q, occ, N = W.shape
    Wv = W.reshape((Q * occ, N), order="F")
    for idx in blocks_to_calculate:
        bp, bq = batch[idx]
        ps, qs = (bp - 1) * width, (bq - 1) * width
        blk = sblas.dgemm(-1.0, Wv[:, ps:ps + width], Wv[:, qs:qs + width], trans_a=1)
        fock[ps:ps + width, qs:qs + width] = blk
        if bp != bq:
            fock[qs:qs + width, ps:ps + width] = blk.T
    if N % n_blocks == 0:
        return
    q0 = N - (N % n_blocks)
    strip = sblas.dgemm(-1.0, Wv, Wv[:, q0:], trans_a=1)  # the remainder strip is never screened (:519-544)
    fock[:, q0:] = strip
    fock[q0:, :] = strip.T


# --------------------------------------------------------------------------------------------------------------
# a4  B formation (DensityFitting.jl:137-152 / ScreenedDF.jl:92-103)
# --------------------------------------------------------------------------------------------------------------
def metric_inverse_cholesky(J_PQ: np.ndarray) -> np.ndarray:
    """potrf!('L') then trtri!('L','N') on the 2-centre metric; only the lower triangle of J_PQ is read.
    Returns Linv (lower triangular, upper triangle zeroed for convenience)."""
    a = np.array(J_PQ, dtype=np.float64, order="F")
    c, info = slapack.dpotrf(a, lower=1, clean=0, overwrite_a=0)
    if info != 0:
        raise np.linalg.LinAlgError("dpotrf info=%d" % info)
    inv, info = slapack.dtrtri(c, lower=1, unitdiag=0)
    if info != 0:
        raise np.linalg.LinAlgError("dtrtri info=%d" % info)
    return np.tril(inv)


def calculate_B(J_PQ: np.ndarray, T: np.ndarray) -> np.ndarray:
    """B = Linv * T  (BLAS.trmm!('L','L','N','N'); NB: L^-1, not the symmetric inverse square root).
    T is (Q, P) column-major (Q fastest)."""
    Linv = metric_inverse_cholesky(J_PQ)
    T = np.asfortranarray(T, dtype=np.float64)
    return sblas.dtrmm(1.0, Linv, T, side=0, lower=1, trans_a=0, diag=0)


def calculate_B_rows(Linv: np.ndarray, T: np.ndarray, q_range: range) -> np.ndarray:
    """Rows q_range of B for one Q shard: B[q_range, :] = Linv[q_range, :] * T — what a shard owns after the
    reference's multi-rank B formation (DensityFitting.jl:164-174 / ScreenedDF.jl:164-186)."""
    return np.asfortranarray(Linv[q_range.start:q_range.stop, :] @ np.asfortranarray(T))


# --------------------------------------------------------------------------------------------------------------
# a2, a3  dense algorithm (DensityFitting.jl:185-224)
# --------------------------------------------------------------------------------------------------------------
def calculate_coulomb_dense(B: np.ndarray, C_occ: np.ndarray, fock: np.ndarray) -> np.ndarray:
    """B: (Q, N, N) 'F'.  density = C_occ*C_occ^T (:193, no factor 2); V = B*vec(density) (:195);
    fock = 2*B^T*V with beta = 0 (:198 — overwrites fock).  Returns V."""
    Q, N, _ = B.shape
    density = sblas.dgemm(1.0, C_occ, C_occ, trans_b=1)
    B2 = B.reshape((Q, N * N), order="F")
    V = sblas.dgemv(1.0, B2, density.reshape(-1, order="F"))
    J = sblas.dgemv(2.0, B2, V, trans=1)
    fock[:, :] = J.reshape((N, N), order="F")
    return V


def calculate_exchange_dense(B: np.ndarray, C_occ: np.ndarray, fock: np.ndarray) -> np.ndarray:
    """W[i,(Q,p)] = sum_q C_occ[q,i]*B[Q,p,q] (:216);  fock -= W'^T W' with W' = reshape(W, (occ*Q, N)) (:219).
    Returns W as (occ, Q*N) 'F'."""
    Q, N, _ = B.shape
    occ = C_occ.shape[1]
    Bm = B.reshape((Q * N, N), order="F")
    W = sblas.dgemm(1.0, C_occ, Bm, trans_a=1, trans_b=1)  # (occ, Q*N)
    W = np.asfortranarray(W)
    Wv = W.reshape((occ * Q, N), order="F")
    fock[:, :] = sblas.dgemm(-1.0, Wv, Wv, beta=1.0, c=np.asfortranarray(fock), trans_a=1)
    return W


# --------------------------------------------------------------------------------------------------------------
# a8-a10  default screened algorithm (ScreenedDF.jl)
# --------------------------------------------------------------------------------------------------------------
def calculate_W_screened(B: np.ndarray, C_occ_T: np.ndarray, sd: ScreeningData) -> np.ndarray:
    """ScreenedDF.jl:242-289.  B: (Q, P) 'F' in the reference's pair order; C_occ_T: (occ, N) (the permutedims copy
    made at :84).  For each p gathers the kept columns of C_occ_T (:265-270) and W[:, :, p] = B[:, s_p : s_p+K_p-1] *
    gathered^T (:271-275).  Returns W = D_tilde (Q, occ, N) 'F'."""
    Q = B.shape[0]
    occ, N = C_occ_T.shape
    W = np.zeros((Q, occ, N), dtype=np.float64, order="F")
    mask = sd.basis_function_screen_matrix
    for p in range(N):
        nz = np.asfortranarray(C_occ_T[:, mask[:, p]])  # (occ, K_p)
        K_p = int(sd.non_screened_p_indices_count[p])
        s = int(sd.sparse_p_start_indices[p]) - 1
        W[:, :, p] = sblas.dgemm(1.0, B[:, s:s + K_p], nz, trans_b=1)
    return W


def calculate_K_lower_diagonal_block_no_screen(W: np.ndarray, fock: np.ndarray, n_blocks: int, width: int, batch):
    """ScreenedDF.jl:548-641 (default: df_exchange_screen = false).  Per lower-triangular block (bp, bq):
    fock[bp, bq] = -W[:, :, bp]^T W[:, :, bq] (alpha = -1, beta = 0 — OVERWRITES fock) mirrored to [bq, bp]
    (:596-601); the N % n_blocks remainder strip is one extra GEMM against all p (:619-640)."""
    Q, occ, N = W.shape
    Wv = W.reshape((Q * occ, N), order="F")
    for (bp, bq) in batch:
        ps, qs = (bp - 1) * width, (bq - 1) * width
        blk = sblas.dgemm(-1.0, Wv[:, ps:ps + width], Wv[:, qs:qs + width], trans_a=1)
        fock[ps:ps + width, qs:qs + width] = blk
        if bp != bq:
            fock[qs:qs + width, ps:ps + width] = blk.T
    if N % n_blocks == 0:
        return
    q0 = N - (N % n_blocks)
    strip = sblas.dgemm(-1.0, Wv, Wv[:, q0:], trans_a=1)  # (N, rem)
    fock[:, q0:] = strip
    fock[q0:, :] = strip.T


def copy_screened_density_to_array(density: np.ndarray, sd: ScreeningData) -> np.ndarray:
    """ScreenedDF.jl:305-316: d[idx(r,c)] = 2*Dm[r,c] (r > c), Dm[r,r] on the diagonal, only r >= c and kept."""
    d = np.zeros(sd.screened_indices_count, dtype=np.float64)
    mask, imap = sd.basis_function_screen_matrix, sd.sparse_pq_index_map
    r, c = np.nonzero(np.tril(mask))
    vals = np.where(r != c, 2.0 * density[r, c], density[r, c])
    d[imap[r, c] - 1] = vals
    return d


def calculate_coulomb_screened(B: np.ndarray, C_occ_T: np.ndarray, sd: ScreeningData, fock: np.ndarray):
    """ScreenedDF.jl:318-378.  density = C_occ_T^T C_occ_T (:320); V += B[:, tail(p)] d[tail(p)] per p (:329-340)
    where tail(p) = sparse indices idx(p,p) .. start(p+1)-1 (the q >= p part of column p); J[tail] += 2 B[:, tail]^T V
    (:346-357); fock[r,c] += J[idx(r,c)] for r >= c, mirrored (:367-378).  Returns (V, Jvec)."""
    occ, N = C_occ_T.shape
    density = sblas.dgemm(1.0, C_occ_T, C_occ_T, trans_a=1)
    d = copy_screened_density_to_array(density, sd)
    imap, starts = sd.sparse_pq_index_map, sd.sparse_p_start_indices
    P = sd.screened_indices_count
    Q = B.shape[0]
    V = np.zeros(Q, dtype=np.float64)
    tails = []
    for p in range(N - 1):
        lo = int(imap[p, p]) - 1
        hi = int(starts[p + 1]) - 1  # exclusive, 0-based
        tails.append((lo, hi))
    tails.append((P - 1, P))
    for lo, hi in tails:
        V = sblas.dgemv(1.0, B[:, lo:hi], d[lo:hi], beta=1.0, y=V)
    Jv = np.zeros(P, dtype=np.float64)
    for lo, hi in tails:
        Jv[lo:hi] = sblas.dgemv(2.0, B[:, lo:hi], V, beta=1.0, y=Jv[lo:hi], trans=1)
    mask = sd.basis_function_screen_matrix
    r, c = np.nonzero(np.tril(mask))
    fock[r, c] += Jv[imap[r, c] - 1]
    fock[c, r] = fock[r, c]
    return V, Jv


# --------------------------------------------------------------------------------------------------------------
# a1  dispatcher (DensityFitting.jl:23-76) — one rank's contribution and the all-rank sum
# --------------------------------------------------------------------------------------------------------------
def fock_build_dense_rank(B_dense: np.ndarray, C_occ: np.ndarray) -> np.ndarray:
    """df_rhf_fock_build_BLAS! per-iteration part (DensityFitting.jl:123-124) for one rank's Q rows."""
    N = B_dense.shape[1]
    fock = np.zeros((N, N), dtype=np.float64, order="F")
    calculate_coulomb_dense(B_dense, np.asfortranarray(C_occ), fock)
    calculate_exchange_dense(B_dense, np.asfortranarray(C_occ), fock)
    return fock


def fock_build_screened_rank(B: np.ndarray, C_occ: np.ndarray, sd: ScreeningData,
                             df_exchange_n_blocks: int = 0, df_screen_exchange: bool = False,
                             two_electron_fock: np.ndarray | None = None) -> np.ndarray:
    """df_rhf_fock_build_screened! per-iteration part (ScreenedDF.jl:84, 130-131) for one rank's Q rows:
    exchange first (overwrites fock), then Coulomb (+=).  `two_electron_fock` is the rank's scf_data.two_electron_fock as
    the iteration finds it (zeros at iteration 1, DensityFitting.jl:45); it only matters with df_screen_exchange, whose
    dropped blocks are never written (calculate_exchange_screened!, :231-235)."""
    C_occ_T = np.asfortranarray(np.asarray(C_occ).T)
    N = C_occ_T.shape[1]
    fock = np.zeros((N, N), dtype=np.float64, order="F") if two_electron_fock is None else \
        np.array(two_electron_fock, dtype=np.float64, order="F", copy=True)
    n_blocks, width, batch = exchange_block_layout(N, df_exchange_n_blocks, 10)
    W = calculate_W_screened(B, C_occ_T, sd)
    if df_screen_exchange:
        todo, _ = exchange_block_screen_matrix(sd.basis_function_screen_matrix, n_blocks, width, batch, True)
        calculate_K_lower_diagonal_block(W, fock, n_blocks, width, batch, todo)
    else:
        calculate_K_lower_diagonal_block_no_screen(W, fock, n_blocks, width, batch)
    calculate_coulomb_screened(B, C_occ_T, sd, fock)
    return fock


def df_rhf_fock_build(B_shards, C: np.ndarray, occ: int, H: np.ndarray, sd: ScreeningData | None = None,
                      contraction_mode: str = "screened", df_exchange_n_blocks: int = 0,
                      df_screen_exchange: bool = False, two_electron_fock: np.ndarray | None = None) -> np.ndarray:
    """df_rhf_fock_build! (DensityFitting.jl:23-76) with the MPI ranks emulated as a list of shards:
    C_occ = C[:, 1:occ] (:49); every rank builds its partial two-electron Fock from its Q rows; rank 0 adds H (:62-65);
    Allreduce SUM (:68-71).  B_shards[r] is (Q_r, N, N) for "dense" or (Q_r, P) for "screened".
    `two_electron_fock`: what scf_data.two_electron_fock holds on EVERY rank when the call starts — the matrix the previous
    call returned, as mutated by the (replicated) SCF loop; None = zeros (iteration 1).  Read only by df_screen_exchange."""
    C_occ = np.asfortranarray(C[:, :occ])
    total = None
    for rank, Bs in enumerate(B_shards):
        if Bs.shape[0] == 0:
            # a rank that owns no auxiliary shell contributes a zero matrix (BLAS calls with M = 0) — except in the blocks
            # df_screen_exchange never writes, which keep the rank's old content
            f = np.zeros_like(H)
            if df_screen_exchange and contraction_mode != "dense" and two_electron_fock is not None:
                n_blocks, width, batch = exchange_block_layout(H.shape[0], df_exchange_n_blocks, 10)
                _, bsm = exchange_block_screen_matrix(sd.basis_function_screen_matrix, n_blocks, width, batch, True)
                for bp in range(n_blocks):
                    for bq in range(n_blocks):
                        if not (bsm[bp, bq] or bsm[bq, bp]):
                            f[bp * width:(bp + 1) * width, bq * width:(bq + 1) * width] = \
                                two_electron_fock[bp * width:(bp + 1) * width, bq * width:(bq + 1) * width]
        elif contraction_mode == "dense":
            f = fock_build_dense_rank(Bs, C_occ)
        else:
            f = fock_build_screened_rank(Bs, C_occ, sd, df_exchange_n_blocks, df_screen_exchange, two_electron_fock)
        if rank == 0:
            f = f + H
        total = f if total is None else total + f
    return np.asfortranarray(total)


# --------------------------------------------------------------------------------------------------------------
# layout helpers shared by tests: dense (Q,N,N) <-> reference screened (Q,P)
# --------------------------------------------------------------------------------------------------------------
def dense_to_screened(B_dense: np.ndarray, sd: ScreeningData) -> np.ndarray:
    """Gather B_dense[:, q, p] into the reference's (Q, P) pair order (sparse index of (q,p) = map[q,p])."""
    Q, N, _ = B_dense.shape
    B2 = B_dense.reshape((Q, N * N), order="F")
    keep = sd.basis_function_screen_matrix.ravel(order="F")
    return np.asfortranarray(B2[:, keep])


def screened_to_dense(B: np.ndarray, sd: ScreeningData) -> np.ndarray:
    Q = B.shape[0]
    N = sd.basis_function_screen_matrix.shape[0]
    B2 = np.zeros((Q, N * N), dtype=np.float64, order="F")
    keep = sd.basis_function_screen_matrix.ravel(order="F")
    B2[:, keep] = B
    return B2.reshape((Q, N, N), order="F")


# --------------------------------------------------------------------------------------------------------------
# SCF.jl:1072-1125 iteration()
# --------------------------------------------------------------------------------------------------------------
def iteration(F: np.ndarray, H: np.ndarray, ortho: np.ndarray, nocc: int):
    """X^T F X (symm!/gemm! :1083-1084), eigen!(Hermitian) (:1086), C = X U (:1095), D[i,j] = 2 sum_occ C C
    (:1106-1108), E_elec = (D.F + D.H)/2 (:1113-1115).  Returns (E_elec, evals, C, D)."""
    Xt = np.asfortranarray(ortho.T)
    Fp = Xt @ F @ ortho
    Fp = 0.5 * (Fp + Fp.T)  # Hermitian() view reads one triangle; symmetrising is equivalent to rounding
    evals, evecs = sla.eigh(Fp)
    C = ortho @ evecs
    Cocc = C[:, :nocc]
    D = 2.0 * (Cocc @ Cocc.T)
    E = 0.5 * (np.vdot(D, F) + np.vdot(D, H))
    return float(E), evals, np.asfortranarray(C), np.asfortranarray(D)


# --------------------------------------------------------------------------------------------------------------
# threaded form of a8-a10, as the reference runs it: Julia tasks over p / over K blocks with BLAS pinned to one
# thread inside (ScreenedDF.jl:244-289, 476-516, 568-616).  Same arithmetic as the serial functions above; used by
# bench.py's cpu_baseline / --impl reference legs so the CPU number is the reference's own parallel decomposition.
# --------------------------------------------------------------------------------------------------------------
def fock_build_screened_rank_threaded(B: np.ndarray, C_occ: np.ndarray, sd: ScreeningData, n_threads: int,
                                      df_exchange_n_blocks: int = 0, timings: dict | None = None,
                                      W_buffer: np.ndarray | None = None) -> np.ndarray:
    """W_buffer: optional (Q, occ, N) column-major array reused between builds — the reference allocates scf_data.D_tilde
    once at iteration 1 (ScreenedDF.jl:110) and overwrites it in every build (gemm with beta = 0)."""
    import time
    from concurrent.futures import ThreadPoolExecutor

    from threadpoolctl import threadpool_limits

    C_occ_T = np.asfortranarray(np.asarray(C_occ).T)
    occ, N = C_occ_T.shape
    Q = B.shape[0]
    fock = np.zeros((N, N), dtype=np.float64, order="F")
    n_blocks, width, batch = exchange_block_layout(N, df_exchange_n_blocks, 10)
    mask = sd.basis_function_screen_matrix
    W = W_buffer if W_buffer is not None else np.zeros((Q, occ, N), dtype=np.float64, order="F")
    assert W.shape == (Q, occ, N) and W.flags.f_contiguous
    Wv = W.reshape((Q * occ, N), order="F")
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=n_threads) as pool:
        with threadpool_limits(limits=1, user_api="blas"):  # BLAS.set_num_threads(1), ScreenedDF.jl:246
            def w_of_p(p):
                nz = np.asfortranarray(C_occ_T[:, mask[:, p]])
                K_p = int(sd.non_screened_p_indices_count[p])
                s = int(sd.sparse_p_start_indices[p]) - 1
                W[:, :, p] = np.dot(B[:, s:s + K_p], nz.T)  # numpy's dgemm call releases the GIL (f2py's does not)
            list(pool.map(w_of_p, range(N)))
            t1 = time.perf_counter()

            def k_block(b):
                bp, bq = b
                ps, qs = (bp - 1) * width, (bq - 1) * width
                blk = -np.dot(Wv[:, ps:ps + width].T, Wv[:, qs:qs + width])
                fock[ps:ps + width, qs:qs + width] = blk
                if bp != bq:
                    fock[qs:qs + width, ps:ps + width] = blk.T
            list(pool.map(k_block, batch))
        if N % n_blocks != 0:  # remainder strip: one multi-threaded GEMM (ScreenedDF.jl:619-640)
            q0 = N - (N % n_blocks)
            strip = sblas.dgemm(-1.0, Wv, Wv[:, q0:], trans_a=1)
            fock[:, q0:] = strip
            fock[q0:, :] = strip.T
    t2 = time.perf_counter()
    calculate_coulomb_screened(B, C_occ_T, sd, fock)  # gemv sweeps with threaded BLAS (ScreenedDF.jl:326-358)
    t3 = time.perf_counter()
    if timings is not None:
        timings.update({"W": t1 - t0, "K": t2 - t1, "J": t3 - t2})
    return fock


def fock_build_screened_slabs(get_rows, Q: int, C_occ: np.ndarray, sd: ScreeningData, n_threads: int, slab: int,
                              timings: dict | None = None) -> np.ndarray:
    """The two-electron Fock matrix of all Q auxiliary rows as the sum of fock_build_screened_rank_threaded over
    contiguous row slabs — the way the reference's MPI ranks each contract their own Q rows and Allreduce the partials
    (DensityFitting.jl:68-71).  get_rows(q0, q1) returns rows [q0, q1) of B as a (q1 - q0, P) column-major array, so a
    tensor too large for one host array (real w50: 7.7 GB) is checked slab by slab."""
    total = None
    for q0 in range(0, Q, slab):
        q1 = min(Q, q0 + slab)
        tm = {}
        f = fock_build_screened_rank_threaded(np.asfortranarray(get_rows(q0, q1)), C_occ, sd, n_threads, 0, tm)
        total = f if total is None else total + f
        if timings is not None:
            for k, v in tm.items():
                timings[k] = timings.get(k, 0.0) + v
    return total
