"""Pins the oracle against the reference's OWN logged run (water_ccpvdz_out.log -> tests/golden/water_ccpvdz_rifit.json,
made by tests/golden/make_water_golden.py): integrals from oracle/gaussian_integrals.py, B = L^-1 (P|mn) and the Fock
build from oracle/df_oracle.py, driven through oracle/scf_replay.py.  What the log holds and what is compared:

  * the 11 printed iteration lines (E to 10 decimals, D_rms to 10 decimals)            -> |dE| <= 2e-8 Eh, |dDrms| <= 1e-7
  * "Total SCF Energy: -75.9911548795 h" (the 12th, unprinted, iteration; SCF.jl:527-546 breaks before @printf)
                                                                                        -> |dE| <= 5e-10 Eh
  * the 25 x 25 matrix dumped before every iteration line (8 significant decimals): it is the two-electron Fock
    matrix 2J - K of that iteration's build                                             -> element-wise <= 1.5e-7
The basis-set digits printed in the log (6 decimals) are the published cc-pVDZ / cc-pVDZ-RIFIT values, so no tolerance
is needed for them.  CPU only."""
import numpy as np
import pytest

from oracle import df_oracle as orc
from oracle import scf_replay as sr
from oracle import water_case as wc


@pytest.fixture(scope="module")
def water():
    g = wc.load_golden()
    p = wc.build_problem(g)
    return g, p


def _replay(p, builder):
    mats = {}
    res = sr.run_scf(p["S"], p["H"], p["E_nuc"], p["nocc"], builder, record=lambda it, k, v: mats.setdefault(it, v))
    return res, mats


def test_integrals_are_normalised_and_metric_is_spd(water):
    g, p = water
    assert p["N"] == g["settings"]["N"] == 25 and p["Q"] == g["settings"]["Naux"] == 96
    assert np.abs(np.diag(p["S"]) - 1.0).max() < 1e-12
    assert np.abs(p["J_PQ"] - p["J_PQ"].T).max() < 1e-12
    assert np.linalg.eigvalsh(p["J_PQ"]).min() > 0
    assert np.abs(p["T3"] - p["T3"].transpose(0, 2, 1)).max() < 1e-13


def test_dense_algorithm_reproduces_the_logged_scf(water):
    g, p = water
    B, H, nocc = p["B"], p["H"], p["nocc"]
    res, mats = _replay(p, lambda C, it: orc.df_rhf_fock_build([B], C, nocc, H, None, "dense"))
    assert res["converged"] and len(res["iterations"]) == 12
    for (it, E, dE, drms), gl in zip(res["iterations"], g["iterations"]):
        assert it == gl["iter"]
        assert abs(E - gl["E"]) <= 2e-8, (it, E, gl["E"])
        assert abs(drms - gl["D_rms"]) <= 1e-7
    assert abs(res["E"] - g["E_converged"]) <= 5e-10
    assert np.abs((mats[1] - H) - np.array(g["matrix_iter1"])).max() <= 1.5e-7
    assert np.abs((mats[11] - H) - np.array(g["matrix_last"])).max() <= 1.5e-7


def test_screened_algorithm_and_q_sharding_match_dense_on_water(water):
    """Default CPU algorithm (ScreenedDF.jl) with the all-true mask water gets at sigma = 1e-5, 1 rank and 3 ranks
    sharded by auxiliary shells (DynamicLoad.jl:160-203): same Fock matrix as the dense algorithm to rounding."""
    g, p = water
    B, H, nocc, N, Q = p["B"], p["H"], p["nocc"], p["N"], p["Q"]
    res, _ = _replay(p, lambda C, it: orc.df_rhf_fock_build([B], C, nocc, H, None, "dense"))
    C = res["C"]
    F_dense = orc.df_rhf_fock_build([B], C, nocc, H, None, "dense")
    sd = orc.ScreeningData(np.ones((N, N), dtype=bool))
    Bs = orc.dense_to_screened(B, sd)
    F_scr = orc.df_rhf_fock_build([Bs], C, nocc, H, sd, "screened")
    assert np.abs(F_scr - F_dense).max() < 1e-11
    ranges = [orc.static_load_rank_indicies(r, 3, p["aux_shell_nbas"])[1] for r in range(3)]
    assert sum(len(r) for r in ranges) == Q
    F_sh = orc.df_rhf_fock_build([np.asfortranarray(Bs[r.start:r.stop]) for r in ranges], C, nocc, H, sd, "screened")
    assert np.abs(F_sh - F_dense).max() < 1e-11
