"""Row a7 with df_exchange_screen = true: libjcdf's option "exchange_block_screen" against the oracle's restatement of
calculate_exchange_block_screen_matrix + calculate_K_lower_diagonal_block (ScreenedDF.jl:385-545) through the C-ABI.

The reference computes K only for the blocks of its df_exchange_n_blocks grid whose sub-block of the Schwarz mask holds a
pair, and it neither computes nor WRITES the others: there scf_data.two_electron_fock keeps what it held — the previous
Fock matrix after the caller's DIIS and damping — plus H on rank 0, summed over the ranks.  So every test here runs a
SEQUENCE of builds over one caller-owned buffer that is mutated between the calls, exactly as SCF.jl:463-508 does, and
compares every matrix of the sequence with the oracle's (1e-10 per element, scaled by the growth of the kept copies)."""
import numpy as np
import pytest

from helpers import banded_mask, jc, orc, random_inputs, reference_layout_rows

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _mutate(F, it):
    """Stand-in for DIIS + damping (SCF.jl:471-508): any in-place change of the array the build returned."""
    F *= 0.75
    F += 0.01 * it * np.eye(F.shape[0])


def _sequence(N, Q, occ, mask, n_blocks, shards, ht_mode, n_builds=3, seed=0, ndev=None, regroup=1):
    ndev = ndev or len(shards)
    B, _, H = random_inputs(N, Q, occ, seed=seed, mask=mask, decay=N / 8.0)
    sd = orc.ScreeningData(mask)
    Bs = [orc.dense_to_screened(np.asfortranarray(B[r.start:r.stop]), sd) for r in shards]
    gd = jc.B200Data(N, Q, occ, ndev)
    gd.set_shards(shards)
    gd.set_option("ht_mode", ht_mode)
    gd.set_option("regroup", regroup)
    gd.set_option("exchange_block_screen", n_blocks)
    gd.set_mask(jc.ScreeningData(mask, N))
    for d, r in enumerate(shards):
        gd.set_B(d, reference_layout_rows(B, r, mask))
    F_gpu = np.zeros((N, N), order="F")  # scf_data.two_electron_fock = zeros at iteration 1 (DensityFitting.jl:45)
    F_orc = None
    errs = []
    for it in range(1, n_builds + 1):
        C_occ = random_inputs(N, 1, occ, seed=seed + 100 * it)[1]
        Cfull = np.zeros((N, N), order="F")
        Cfull[:, :occ] = C_occ
        F_orc = orc.df_rhf_fock_build(Bs, Cfull, occ, H, sd, "screened", df_exchange_n_blocks=n_blocks,
                                      df_screen_exchange=True, two_electron_fock=F_orc)
        out, _ = gd.fock_build(C_occ, H, F_gpu)
        assert out is F_gpu
        assert np.array_equal(F_gpu, F_gpu.T)
        errs.append(np.abs(F_gpu - F_orc).max() / max(1.0, np.abs(F_orc).max()))
        _mutate(F_gpu, it)
        _mutate(F_orc, it)
    return gd, errs, (B, H, sd, Bs)


@pytest.mark.parametrize("ht_mode", [1, 2])
@pytest.mark.parametrize("N,Q,occ,hw,n_blocks", [(120, 16, 7, 18, 10),    # width 12, square grid covers N
                                                 (203, 40, 13, 30, 10),   # width 20 + a never-screened strip of 3
                                                 (300, 48, 21, 40, 20),   # width 15; whole 128 x 128 tiles drop out
                                                 (150, 24, 9, 25, 7),     # width 21, strip of 3
                                                 (470, 24, 11, 45, 47)])  # width 10; skipped tiles, dead 64-row halves, clipped last tile row
def test_block_screened_sequence_matches_the_oracle(N, Q, occ, hw, n_blocks, ht_mode):
    mask = banded_mask(N, hw, seed=N)
    gd, errs, _ = _sequence(N, Q, occ, mask, n_blocks, [range(0, Q)], ht_mode, seed=N + 3)
    li = gd.layout()
    nb, width, batch = orc.exchange_block_layout(N, n_blocks, 10)
    todo, _ = orc.exchange_block_screen_matrix(mask, nb, width, batch, True)
    assert (li.exchange_blocks_total, li.exchange_blocks_kept, li.exchange_block_width) == (len(batch), len(todo), width)
    assert 0 < len(todo) < len(batch)
    if N == 300:
        assert li.exchange_tiles_total == 6 and li.exchange_tiles_skipped >= 1  # tile (2, 0): |p - q| > 128 > band + block
    if N in (300, 470):  # 64-row halves of the remaining tiles whose warps compute nothing (identity function order: 1 / 2 of them)
        assert li.exchange_half_tiles_skipped >= 1
    assert max(errs) < TOL, errs
    gd.close()


def test_dropped_blocks_carry_the_callers_buffer_and_kept_blocks_equal_the_default_build():
    N, Q, occ, n_blocks = 203, 40, 13, 10
    mask = banded_mask(N, 30, seed=N)
    B, C_occ, H = random_inputs(N, Q, occ, seed=77, mask=mask, decay=N / 8.0)
    nb, width, batch = orc.exchange_block_layout(N, n_blocks, 10)
    _, bsm = orc.exchange_block_screen_matrix(mask, nb, width, batch, True)
    dropped = np.zeros((N, N), dtype=bool)
    dropped[:nb * width, :nb * width] = ~np.kron(bsm | bsm.T, np.ones((width, width), dtype=bool))
    res = {}
    for opt in (0, n_blocks):
        gd = jc.B200Data(N, Q, occ, 1)
        gd.set_option("exchange_block_screen", opt)
        gd.set_mask(jc.ScreeningData(mask, N))
        gd.set_B(0, reference_layout_rows(B, range(0, Q), mask))
        old = np.asfortranarray(np.arange(N * N, dtype=np.float64).reshape(N, N))
        old = np.asfortranarray(old + old.T)
        F = old.copy(order="F")
        gd.fock_build(C_occ, H, F)
        res[opt] = (F.copy(), old)
        # a second build over the SAME buffer content gives the same matrix again (bit for bit): nothing is kept inside the handle
        F2 = old.copy(order="F")
        gd.fock_build(C_occ, H, F2)
        assert np.array_equal(F, F2)
        gd.close()
    F0, _ = res[0]
    Fs, old = res[n_blocks]
    assert np.abs(Fs - F0)[~dropped].max() < 1e-12              # computed blocks: the default build's values
    assert np.array_equal(Fs[dropped], (old + H)[dropped])     # dropped blocks: old content + H, nothing else
    assert not np.array_equal(F0[dropped], Fs[dropped])


def test_below_100_functions_there_is_one_block_and_a_null_mask_ignores_the_option():
    """ScreenedDF.jl:392-394 (N < 100: K_block_width = N, one block) and DensityFitting.jl:185-224 (dense: no blocks)."""
    N, Q, occ = 70, 40, 9
    mask = banded_mask(N, 12, seed=1)
    gd, errs, (B, H, sd, Bs) = _sequence(N, Q, occ, mask, 10, [range(0, Q)], 0, seed=5)
    li = gd.layout()
    assert (li.exchange_blocks_total, li.exchange_blocks_kept, li.exchange_block_width) == (1, 1, N)
    assert max(errs) < TOL
    gd.close()
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_option("exchange_block_screen", 10)
    gd.set_mask(jc.ScreeningData(None, N))
    assert gd.layout().exchange_blocks_total == 0
    gd.close()
    gd = jc.B200Data(N, Q, occ, 1)
    gd.set_mask(jc.ScreeningData(mask, N))
    with pytest.raises(jc.JCDFError):
        gd.set_option("exchange_block_screen", 10)  # too late: the block grid is laid out by jcdf_set_mask
    gd.close()


@pytest.mark.parametrize("ndev", [2, 3])
def test_block_screened_sequence_on_several_devices(ndev):
    """Every shard acts as one rank of the reference: each keeps a copy of the old matrix in the dropped blocks and the
    all-reduce sums them (n_ranks * old + H), DensityFitting.jl:62-71."""
    if jc.device_count() < ndev:
        pytest.skip("needs %d GPUs" % ndev)
    N, Q, occ, n_blocks = 203, 41, 13, 10
    mask = banded_mask(N, 30, seed=N)
    shards = jc.balanced_function_shards(Q, ndev)
    gd, errs, _ = _sequence(N, Q, occ, mask, n_blocks, shards, 0, seed=11)
    assert max(errs) < TOL, errs
    gd.close()


def test_seam_honours_df_exchange_screen_over_three_iterations():
    """df_rhf_fock_build_B200 with SCFOptions(df_screen_exchange = True): scf_data.two_electron_fock is the buffer the
    library reads and writes, the caller mutates it between the iterations, and the timing object books the block counts
    calculate_exchange_block_screen_matrix books (ScreenedDF.jl:409-410, 454)."""
    N, Q, occ = 160, 30, 11
    mask = banded_mask(N, 22, seed=8)
    B, _, H = random_inputs(N, Q, occ, seed=21, mask=mask, decay=20.0)
    sd = orc.ScreeningData(mask)
    Bs = [orc.dense_to_screened(B, sd)]

    class Src(jc.IntegralSource):
        def __init__(self):
            self.N, self.Q, self.n_electrons = N, Q, 2 * occ

        def screening_matrix(self, sigma):
            return mask

        def B_rows(self, q_range, sdata):
            return reference_layout_rows(B, q_range, mask)

    opts = jc.SCFOptions(df_screen_exchange=True, df_exchange_n_blocks=0, df_use_adaptive=False)
    scf_data, timing = jc.SCFData(), jc.JCTiming()
    F_orc = None
    for it in (1, 2, 3):
        C_occ = random_inputs(N, 1, occ, seed=300 + it)[1]
        Cfull = np.zeros((N, N), order="F")
        Cfull[:, :occ] = C_occ
        F = jc.df_rhf_fock_build_B200(scf_data, Src(), C_occ, it, opts, H, timing)
        assert F is scf_data.two_electron_fock
        F_orc = orc.df_rhf_fock_build(Bs, Cfull, occ, H, sd, "screened", df_screen_exchange=True, two_electron_fock=F_orc)
        assert np.abs(F - F_orc).max() < TOL * max(1.0, np.abs(F_orc).max())
        _mutate(F, it)
        _mutate(F_orc, it)
    nb, width, batch = orc.exchange_block_layout(N, 0, 10)
    todo, _ = orc.exchange_block_screen_matrix(mask, nb, width, batch, True)
    assert timing.non_timing_data["total_exchange_blocks"] == str(len(batch)) == "55"
    assert timing.non_timing_data["unscreened_exchange_blocks"] == str(len(todo))
    assert timing.non_timing_data["df_exchange_n_blocks"] == "10"
    scf_data.gpu_data.close()


def test_resident_scf_loop_with_block_screened_exchange_on_a_spread_water_tetramer():
    """Whole path from the geometry on (device integrals, device mask, resident SCF loop: DIIS and damping act on the very
    array the build reads back) against the oracle's replay of SCF.jl with the same keyword.  Waters 1, 2, 46 and 48 of the
    w50 cluster in cc-pVDZ (N = 100, one water per K block at df_exchange_n_blocks = 4): the O-O distances of 8-10 A drop
    3 of the 10 blocks.  With the keyword on, the reference's iteration sequence is no longer a convergent one on one rank
    (the dropped blocks gain another H every iteration), so the first iterations are compared, energy by energy, to the
    north star's 1e-8 Eh, without DIIS (8 iterations) and with it (6; measured differences 6e-12 .. 8e-11 Eh)."""
    import json
    import os
    from juliachem_jl_b200 import rhf
    from oracle import gaussian_integrals as gi, scf_replay as sr
    geo = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "w50_geometry.json")))
    g, sym = [], []
    for m in (0, 1, 45, 47):
        g += geo["geometry"][9 * m:9 * m + 9]
        sym += geo["symbols"][3 * m:3 * m + 3]
    mol = rhf.run_basis({"geometry": g, "symbols": sym, "molecular_charge": 0}, {"basis": "cc-pVDZ"})
    lib = rhf.load_basis_library()
    prim, aux = [], []
    for ia, s in enumerate(mol["symbols"]):
        prim += [gi.Shell(sh["l"], np.array(sh["exps"]), np.array(sh["coefs"]), mol["centers"][ia], ia) for sh in lib[s]["basis"]]
        aux += [gi.Shell(sh["l"], np.array(sh["exps"]), np.array(sh["coefs"]), mol["centers"][ia], ia) for sh in lib[s]["auxiliary_basis"]]
    S, T = gi.overlap_kinetic(prim)
    H = np.asfortranarray(T + gi.nuclear_attraction(prim, mol["charges"], mol["centers"]))
    E_nuc = gi.nuclear_repulsion(mol["charges"], mol["centers"])
    J = gi.two_center_integrals(aux)
    T3 = gi.three_center_integrals(aux, prim)
    Q, N = T3.shape[0], T3.shape[1]
    thr = orc.schwarz_threshold(1e-5, float(np.abs(np.diag(J)).max()))
    pos, _ = gi.shell_offsets(prim)
    _, fn_mask = orc.schwarz_function_mask(gi.schwarz_blocks(prim), pos, [s.nbas for s in prim], thr)
    sd = orc.ScreeningData(fn_mask)
    nb, width, batch = orc.exchange_block_layout(N, 4, 10)
    todo, _ = orc.exchange_block_screen_matrix(fn_mask, nb, width, batch, True)
    assert (N, nb, width, len(batch), len(todo)) == (100, 4, 25, 10, 7)
    Bs = orc.calculate_B(np.tril(J), np.asfortranarray(T3.reshape(Q, N * N))[:, fn_mask.ravel(order="F")])
    nocc = mol["n_electrons"] // 2

    def oracle_build(C, it, F_prev):
        return orc.df_rhf_fock_build([Bs], C, nocc, H, sd, "screened", df_exchange_n_blocks=4, df_screen_exchange=True,
                                     two_electron_fock=F_prev)

    for ndiis, niter, tol in ((0, 8, 1e-8), (10, 6, 1e-8)):
        ref = sr.run_scf(S, H, E_nuc, nocc, oracle_build, ndiis=ndiis, max_iterations=niter, pass_previous_fock=True)
        res = rhf.run_energy(mol, {"df_sigma": 1e-5, "df_exchange_screen": True, "df_exchange_n_blocks": 4, "ndiis": ndiis,
                                   "niter": niter})
        assert res["pairs_kept"] == int(fn_mask.sum())
        li = res["layout"]
        assert (li.exchange_blocks_total, li.exchange_blocks_kept, li.exchange_block_width) == (10, 7, 25)
        assert len(res["Iterations"]) == len(ref["iterations"]) == niter
        diffs = [abs(E - E_ref) for (_, E, _, _), (_, E_ref, _, _) in zip(res["Iterations"], ref["iterations"])]
        print("block-screened SCF, ndiis=%d: max |E_gpu - E_oracle| per iteration = %s" % (ndiis, ["%.1e" % d for d in diffs]))
        for (it, E, _, _), (_, E_ref, _, _) in zip(res["Iterations"], ref["iterations"]):
            assert abs(E - E_ref) < tol, (ndiis, it, E, E_ref)
    # and the keyword changes the run: without it the same tetramer gives another energy at the same iteration
    plain = rhf.run_energy(mol, {"df_sigma": 1e-5, "ndiis": 10, "niter": 6})
    assert abs(plain["Iterations"][-1][1] - res["Iterations"][-1][1]) > 1e-6
