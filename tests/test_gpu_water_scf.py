"""Real-molecule parity on the GPU: the reference's logged water cc-pVDZ / cc-pVDZ-RIFIT DF-RHF run
(tests/golden/water_ccpvdz_rifit.json) replayed with libjcdf doing every Fock build — from raw (P|Q) and (P|mn):
on-device potrf + trtri + B formation, then J/K per iteration — through the host-side mirror of df_rhf_fock_build!.

Tolerances (BASELINE.json north_star): 1e-10 per Fock element against the oracle on the same inputs, 1e-8 Eh on the
converged energy; the logged energies (10 decimals) are matched to 2e-8."""
import numpy as np
import pytest

from helpers import jc, orc
from oracle import scf_replay as sr
from oracle import water_case as wc

pytestmark = pytest.mark.gpu


class WaterIntegrals(jc.IntegralSource):
    """Plays the JERI engines + basis_sets arguments: hands the raw integrals to the B200 path in reference layout."""

    def __init__(self, p, sharded_by_shell=True):
        self.p, self.N, self.Q, self.n_electrons = p, p["N"], p["Q"], 2 * p["nocc"]
        self.by_shell = sharded_by_shell

    def aux_shell_nbas(self):
        return self.p["aux_shell_nbas"] if self.by_shell else None

    def two_center_integrals(self):
        return np.asfortranarray(np.tril(self.p["J_PQ"]))  # lower triangle only, like TwoCenterIntegrals.jl:39-41

    def three_center_integrals(self, q_range, sd):
        T3 = self.p["T3"][q_range.start:q_range.stop]
        return np.asfortranarray(T3.reshape(len(q_range), self.N * self.N))


@pytest.fixture(scope="module")
def water():
    g = wc.load_golden()
    return g, wc.build_problem(g)


@pytest.mark.parametrize("ndev,device_eigh", [(1, False), (1, True), (2, False)])
def test_water_scf_through_libjcdf(water, ndev, device_eigh):
    g, p = water
    if jc.device_count() < ndev:
        pytest.skip("needs %d GPUs" % ndev)
    H, nocc, B = p["H"], p["nocc"], p["B"]
    scf_data, timing = jc.SCFData(), jc.JCTiming()
    opts = jc.SCFOptions(num_devices=ndev)
    src = WaterIntegrals(p)
    worst = [0.0]

    def gpu_fock(C, it):
        F = jc.df_rhf_fock_build_B200(scf_data, src, np.asfortranarray(C[:, :nocc]), it, opts, H, timing)
        F_ref = orc.df_rhf_fock_build([B], C, nocc, H, None, "dense")
        worst[0] = max(worst[0], float(np.abs(F - F_ref).max()))
        return F.copy()  # the caller mutates F (DIIS), SCF.jl:472-508

    it_fn = None
    if device_eigh:
        def it_fn(F, Hc, X, no):
            ev, C = scf_data.gpu_data.eigh(F, X) if scf_data.gpu_data is not None else (None, None)
            if C is None:  # hcore guess happens before the first build created the handle
                return orc.iteration(F, Hc, X, no)
            D = 2.0 * C[:, :no] @ C[:, :no].T
            return float(0.5 * (np.vdot(D, F) + np.vdot(D, Hc))), ev, C, D

    res = sr.run_scf(p["S"], H, p["E_nuc"], nocc, gpu_fock, iteration_fn=it_fn)
    ref = sr.run_scf(p["S"], H, p["E_nuc"], nocc, lambda C, it: orc.df_rhf_fock_build([B], C, nocc, H, None, "dense"))
    assert worst[0] < 1e-10, worst[0]
    assert res["converged"] and len(res["iterations"]) == len(ref["iterations"]) == 12
    assert abs(res["E"] - ref["E"]) < 1e-8
    assert abs(res["E"] - g["E_converged"]) < 1e-9
    for (it, E, dE, drms), gl in zip(res["iterations"], g["iterations"]):
        assert abs(E - gl["E"]) <= 2e-8
    for name in ("W_time", "K_time", "V_time", "J_time", "fock_time"):
        assert jc.JCTiming.key(name, 3) in timing.timings
    scf_data.gpu_data.close()


@pytest.mark.parametrize("ndev", [1, 2])
def test_water_scf_resident_on_device(water, ndev):
    """the whole SCF loop (Fock build, DIIS, damping, iteration()) through jcdf_scf_init/step with F, C, D
    resident on the GPU.  Every iteration's energy against the reference's logged run (<= 2e-8: the log prints 10
    decimals of a run made with another BLAS) and against the oracle's replay (<= 1e-8 Eh), same iteration count."""
    g, p = water
    if jc.device_count() < ndev:
        pytest.skip("needs %d GPUs" % ndev)
    H, nocc, B = p["H"], p["nocc"], p["B"]
    N, Q = p["N"], p["Q"]
    gd = jc.B200Data(N, Q, nocc, ndev)
    gd.set_mask(jc.ScreeningData(None, N))
    gd.upload_metric(np.asfortranarray(np.tril(p["J_PQ"])))
    li = [jc.balanced_function_shards(Q, ndev)[d] for d in range(ndev)]
    for d, r in enumerate(li):
        gd.upload_three_center(d, np.asfortranarray(p["T3"][r.start:r.stop].reshape(len(r), N * N)))
    gd.factor_and_form_B()
    ref = sr.run_scf(p["S"], H, p["E_nuc"], nocc, lambda C, it: orc.df_rhf_fock_build([B], C, nocc, H, None, "dense"))
    E_guess = gd.scf_init(p["S"], H, p["E_nuc"], ndiis=10)
    X = gd.scf_get("X")
    assert np.abs(X - sr.orthogonalizer(p["S"])).max() < 1e-11
    E0, _, _, _ = orc.iteration(np.array(H), H, sr.orthogonalizer(p["S"]), nocc)
    assert abs(E_guess - (E0 + p["E_nuc"])) < 1e-9
    log = []
    for it in range(1, 51):
        r = gd.scf_step()
        log.append(r)
        if abs(r["dE"]) <= 1e-6 and r["D_rms"] <= 1e-6:
            break
    assert len(log) == len(ref["iterations"]) == 12
    for r, (it, E, dE, drms), gl in zip(log, ref["iterations"], g["iterations"]):
        assert abs(r["E"] - E) < 1e-8, (it, r["E"], E)
        assert abs(r["E"] - gl["E"]) <= 2e-8
        assert abs(r["D_rms"] - drms) < 1e-7
    assert abs(log[-1]["E"] - g["E_converged"]) < 1e-9
    # the resident matrices are the Dict rhf_kernel returns (SCF.jl:251-259)
    D, C, F, ev = gd.scf_get("D"), gd.scf_get("C"), gd.scf_get("F"), gd.scf_get("evals")
    assert np.abs(D - ref["D"]).max() < 1e-7
    assert np.abs(ev - ref["evals"]).max() < 1e-7
    assert np.abs(2.0 * C[:, :nocc] @ C[:, :nocc].T - D).max() < 1e-12
    assert np.array_equal(F, F.T) or np.abs(F - F.T).max() < 1e-12
    gd.close()


def test_water_scf_run_whole_loop(water):
    """jcdf_scf_run: the loop itself inside the library, convergence test of SCF.jl:527-530."""
    g, p = water
    N, Q, nocc = p["N"], p["Q"], p["nocc"]
    gd = jc.B200Data(N, Q, nocc, 1)
    gd.set_mask(jc.ScreeningData(None, N))
    B2 = np.asfortranarray(p["B"].reshape((Q, N * N), order="F"))
    gd.set_B(0, B2)
    gd.scf_init(p["S"], p["H"], p["E_nuc"], ndiis=10)
    res = gd.scf_run(50, 1e-6, 1e-6)
    assert res["converged"] and len(res["iterations"]) == 12
    assert abs(res["E"] - g["E_converged"]) < 1e-9
    with pytest.raises(jc.JCDFError):
        gd.scf_get("nonsense")
    gd.close()


class WaterBasisOnly(jc.IntegralSource):
    """The device-integral route of the seam (julia/B200DF.jl: setup_device_integrals_B200!): only the basis sets cross
    the boundary."""

    def __init__(self, g, p):
        from juliachem_jl_b200 import rhf
        xyz = np.array([a["xyz_bohr"] for a in g["geometry"]]) / rhf.ANGSTROM_TO_BOHR
        mol = rhf.run_basis({"geometry": xyz.ravel().tolist(), "symbols": [a["symbol"] for a in g["geometry"]],
                             "molecular_charge": 0}, {"basis": "cc-pVDZ"})
        self.mol, self.N, self.Q, self.n_electrons = mol, p["N"], p["Q"], mol["n_electrons"]

    def aux_shell_nbas(self):
        return self.mol["auxiliary"].nbas

    def device_basis(self):
        return self.mol["primary"].args(), self.mol["auxiliary"].args()


@pytest.mark.parametrize("force_dense", [False, True])
def test_seam_with_device_integrals_and_reference_options(water, force_dense):
    g, p = water
    H, nocc, B = p["H"], p["nocc"], p["B"]
    src = WaterBasisOnly(g, p)
    scf_data, timing = jc.SCFData(), jc.JCTiming()
    opts = jc.SCFOptions(num_devices=1, df_force_dense=force_dense, df_use_adaptive=force_dense, df_exchange_n_blocks=3)
    X = sr.orthogonalizer(p["S"])
    _, _, C, _ = orc.iteration(np.array(H), H, X, nocc)
    F = jc.df_rhf_fock_build_B200(scf_data, src, np.asfortranarray(C[:, :nocc]), 1, opts, H, timing)
    F_ref = orc.df_rhf_fock_build([B], C, nocc, H, None, "dense")
    assert np.abs(F - F_ref).max() < 1e-10
    assert (scf_data.screening_data.basis_function_screen_matrix is None) == force_dense
    # every stage slot lands on a key the reference's JCTiming knows (JCTiming.jl:50-107)
    assert jc.JCTiming.key("GPU_1_non_zero_coeff_time", 1) in timing.timings
    scf_data.gpu_data.close()
    # df_exchange_screen on a molecule below 100 functions: one K block (ScreenedDF.jl:392-394), nothing dropped, same matrix
    sd2, t2 = jc.SCFData(), jc.JCTiming()
    F2 = jc.df_rhf_fock_build_B200(sd2, src, np.asfortranarray(C[:, :nocc]), 1,
                                   jc.SCFOptions(df_screen_exchange=True, df_use_adaptive=False), H, t2)
    assert np.abs(F2 - F_ref).max() < 1e-10
    assert t2.non_timing_data["total_exchange_blocks"] == "1" and t2.non_timing_data["unscreened_exchange_blocks"] == "1"
    assert t2.non_timing_data["df_exchange_n_blocks"] == "1"
    sd2.gpu_data.close()
