"""GPU integral front-end against the oracle's McMurchie-Davidson integrals — the ones that regenerate
the reference's logged water run (tests/test_oracle_water_golden.py) — through the C-ABI:

  jcdf_set_basis / jcdf_compute_two_center / jcdf_compute_schwarz_mask / jcdf_compute_three_center / jcdf_get_B

Tolerance: 1e-12 absolute per integral (values are O(1); the two codes share the scheme but not the Boys evaluation,
recursion layout or summation order), masks and pair counts exact, converged water energy 1e-8 Eh vs the golden log."""
import numpy as np
import pytest

from helpers import jc, orc
from oracle import gaussian_integrals as gi
from oracle import scf_replay as sr
from oracle import water_case as wc

pytestmark = pytest.mark.gpu

TOL = 1e-12


def flat_basis(shells):
    am = [s.l for s in shells]
    nprim = [len(s.exps) for s in shells]
    cen = np.array([s.center for s in shells], dtype=np.float64)
    ex = np.concatenate([np.asarray(s.exps, dtype=np.float64) for s in shells])
    co = np.concatenate([np.asarray(s.coefs, dtype=np.float64) for s in shells])
    return am, nprim, cen, ex, co


def make_handle(prim, aux, nocc=1, ndev=1):
    N, Q = sum(s.nbas for s in prim), sum(s.nbas for s in aux)
    gd = jc.B200Data(N, Q, nocc, ndev)
    gd.set_basis(0, *flat_basis(prim))
    gd.set_basis(1, *flat_basis(aux))
    return gd, N, Q


@pytest.fixture(scope="module")
def water():
    g = wc.load_golden()
    prim = wc.shells_from(g["basis"], g["geometry"])
    aux = wc.shells_from(g["auxiliary_basis"], g["geometry"])
    return g, prim, aux, wc.build_problem(g)


def high_l_shells():
    """Three centres carrying every angular momentum the kernels admit: s..f primary (contracted), s..g auxiliary."""
    cen = [np.array([0.0, 0.0, 0.0]), np.array([1.1, -0.4, 0.7]), np.array([-0.6, 1.3, 0.2])]
    prim = [gi.Shell(0, np.array([3.1, 0.9, 0.3]), np.array([0.2, 0.5, 0.4]), cen[0]),
            gi.Shell(1, np.array([1.7, 0.45]), np.array([0.6, 0.5]), cen[0]),
            gi.Shell(2, np.array([1.2]), np.array([1.0]), cen[1]),
            gi.Shell(3, np.array([0.8]), np.array([1.0]), cen[1]),
            gi.Shell(3, np.array([1.4, 0.5]), np.array([0.3, 0.8]), cen[2]),
            gi.Shell(1, np.array([0.35]), np.array([1.0]), cen[2]),
            gi.Shell(0, np.array([0.15]), np.array([1.0]), cen[1])]
    aux = [gi.Shell(l, np.array([e]), np.array([1.0]), cen[k % 3]) for k, (l, e) in enumerate(
        [(0, 2.5), (0, 0.4), (1, 1.9), (1, 0.5), (2, 1.3), (2, 0.45), (3, 1.1), (3, 0.6), (4, 0.9), (4, 1.6)])]
    aux.append(gi.Shell(2, np.array([2.2, 0.7]), np.array([0.4, 0.7]), cen[1]))  # a contracted auxiliary shell
    return prim, aux


def reference_columns(T3):
    Q, N, _ = T3.shape
    return np.asfortranarray(T3.reshape(Q, N * N))


def test_two_center_water(water):
    g, prim, aux, p = water
    gd, N, Q = make_handle(prim, aux)
    J = gd.compute_two_center()
    assert np.abs(J - p["J_PQ"]).max() < TOL
    assert np.array_equal(J, J.T)
    gd.close()


def test_three_center_water_unscreened(water):
    g, prim, aux, p = water
    gd, N, Q = make_handle(prim, aux)
    gd.compute_two_center(False)
    mask, P = gd.compute_schwarz_mask(0.0)
    assert P == N * N and mask.all()
    gd.compute_three_center()
    T = gd.get_B(0, Q, P)
    assert np.abs(T - reference_columns(p["T3"])).max() < TOL
    gd.close()


@pytest.mark.parametrize("ndev", [1, 2])
def test_high_angular_momentum_all_integrals(ndev):
    """f primary / g auxiliary functions, contracted shells, ragged Q shards that cut through shells."""
    if jc.device_count() < ndev:
        pytest.skip("needs %d GPUs" % ndev)
    prim, aux = high_l_shells()
    gd, N, Q = make_handle(prim, aux, ndev=ndev)
    J_ref = gi.two_center_integrals(aux)
    T_ref = reference_columns(gi.three_center_integrals(aux, prim))
    J = gd.compute_two_center()
    assert np.abs(J - J_ref).max() < TOL
    if ndev == 2:
        cut = Q // 2 + 1  # inside a shell
        shards = [range(0, cut), range(cut, Q)]
        gd.set_shards(shards)
    else:
        shards = [range(0, Q)]
    gd.compute_schwarz_mask(0.0)
    gd.compute_three_center()
    for d, r in enumerate(shards):
        T = gd.get_B(d, len(r), N * N)
        assert np.abs(T - T_ref[r.start:r.stop]).max() < TOL, d
    # B = L^-1 T on the device from the device-resident metric and integrals
    gd.factor_and_form_B()
    B_ref = orc.calculate_B(np.tril(J_ref), T_ref)
    for d, r in enumerate(shards):
        B = gd.get_B(d, len(r), N * N)
        assert np.abs(B - B_ref[r.start:r.stop]).max() < 1e-9 * max(1.0, np.abs(B_ref).max()), d
    # the exchange-free route: every device factors the metric itself and recomputes the rows it needs
    gd.compute_B()
    for d, r in enumerate(shards):
        B = gd.get_B(d, len(r), N * N)
        assert np.abs(B - B_ref[r.start:r.stop]).max() < 1e-9 * max(1.0, np.abs(B_ref).max()), d
    gd.close()


@pytest.mark.parametrize("case,sigma", [("water", 1e-5), ("water", 3e-1), ("water", 1.5), ("high_l", 4e-1), ("stretched", 1e-5)])
def test_schwarz_mask_matches_reference_rule(water, case, sigma):
    """Shell-pair test on the SUM of the (ab|ab) block, function-pair test on its diagonal, compaction outer-p inner-q
    (SchwarzScreening.jl:22-81): masks, pair count and the integrals of the kept pairs."""
    g, prim, aux, p = water
    if case == "high_l":
        prim, aux = high_l_shells()
    elif case == "stretched":  # two waters 14 bohr apart: whole shell pairs fall under the default threshold
        far = [gi.Shell(s.l, s.exps, s.coefs, s.center + np.array([14.0, 0.0, 0.0]), s.atom + 3) for s in prim]
        prim = prim + far
        aux = aux + [gi.Shell(s.l, s.exps, s.coefs, s.center + np.array([14.0, 0.0, 0.0]), s.atom + 3) for s in aux]
    gd, N, Q = make_handle(prim, aux)
    J = gd.compute_two_center()
    thr = orc.schwarz_threshold(sigma, float(np.abs(np.diag(J)).max()))
    pos, _ = gi.shell_offsets(prim)
    shell_mask, fn_mask = orc.schwarz_function_mask(gi.schwarz_blocks(prim), pos, [s.nbas for s in prim], thr)
    mask, P = gd.compute_schwarz_mask(sigma)
    assert np.array_equal(mask, fn_mask)
    assert P == int(fn_mask.sum())
    if case != "water" or sigma > 1e-3:
        assert P < N * N  # the case really screens something
    gd.compute_three_center()
    T = gd.get_B(0, Q, P)
    T_ref = reference_columns(gi.three_center_integrals(aux, prim))[:, fn_mask.ravel(order="F")]
    assert np.abs(T - T_ref).max() < TOL
    assert gd.layout().pairs_kept == P
    gd.close()


def test_water_scf_from_basis_set(water):
    """Everything on the device from the basis set on: (P|Q), Schwarz mask (df_sigma default 1e-5), (P|mn), potrf/trtri,
    B, and the resident SCF loop; energy against the reference's logged run."""
    g, prim, aux, p = water
    gd, N, Q = make_handle(prim, aux, nocc=p["nocc"])
    gd.compute_two_center(False)
    gd.compute_schwarz_mask(1e-5)
    gd.compute_three_center()
    gd.factor_and_form_B()
    gd.scf_init(p["S"], p["H"], p["E_nuc"], ndiis=10)
    res = gd.scf_run(50, 1e-6, 1e-6)
    assert res["converged"] and len(res["iterations"]) == 12
    assert abs(res["E"] - g["E_converged"]) < 1e-8
    for (it, E, dE, drms), gl in zip(res["iterations"], g["iterations"]):
        assert abs(E - gl["E"]) <= 2e-8
    gd.close()


def test_basis_errors(water):
    g, prim, aux, p = water
    gd, N, Q = make_handle(prim, aux)
    am, nprim, cen, ex, co = flat_basis(prim)
    with pytest.raises(jc.JCDFError):
        gd.set_basis(0, am[:-1], nprim[:-1], cen[:-1], ex[:-1], co[:-1])  # functions do not add up to N
    am_bad = list(am)
    am_bad[0] = 5
    with pytest.raises(jc.JCDFError):
        gd.set_basis(0, am_bad, nprim, cen, ex, co)
    gd2 = jc.B200Data(N, Q, 1, 1)
    with pytest.raises(jc.JCDFError):
        gd2.compute_two_center()  # no basis yet
    gd.close()
    gd2.close()


def test_one_electron_integrals(water):
    g, prim, aux, p = water
    for shells, charges, centers in [(prim, g["charges"], [a["xyz_bohr"] for a in g["geometry"]]),
                                     (high_l_shells()[0], [3.0, 1.0, 8.0], [[0.0, 0.0, 0.0], [1.1, -0.4, 0.7], [0.2, 0.9, -1.5]])]:
        N = sum(s.nbas for s in shells)
        gd = jc.B200Data(N, 1, 1, 1)
        gd.set_basis(0, *flat_basis(shells))
        S, T, V, E_nuc = gd.compute_one_electron(charges, centers)
        S_ref, T_ref = gi.overlap_kinetic(shells)
        V_ref = gi.nuclear_attraction(shells, charges, centers)
        assert np.abs(S - S_ref).max() < TOL
        assert np.abs(T - T_ref).max() < 1e-11
        assert np.abs(V - V_ref).max() < 1e-11
        assert abs(E_nuc - gi.nuclear_repulsion(charges, centers)) < 1e-12
        assert np.abs(np.diag(S) - 1.0).max() < 1e-12  # every Cartesian component is unit-normalised
        gd.close()


def test_energy_run_water_json_contract(water):
    """The public entry of the B200 path: molecule/model objects of the JSON contract in, energy Dict out
    (JCBasis.run + JCRHF.Energy.run).  Logged reference energy -75.9911548795."""
    from juliachem_jl_b200 import rhf
    g, prim, aux, p = water
    geo_ang = (np.array([a["xyz_bohr"] for a in g["geometry"]]) / rhf.ANGSTROM_TO_BOHR).ravel().tolist()
    molecule = {"geometry": geo_ang, "symbols": ["O", "H", "H"], "molecular_charge": 0}
    model = {"method": "RHF", "basis": "cc-pVDZ", "auxiliary_basis": "cc-pVDZ-RIFIT"}
    mol = rhf.run_basis(molecule, model)
    assert mol["n_electrons"] == 10 and mol["primary"].n_functions == 25 and mol["auxiliary"].n_functions == 96
    res = rhf.run_energy(mol, {"niter": 50, "ndiis": 10, "dele": 1e-6, "rmsd": 1e-6})
    assert res["Converged?"] and len(res["Iterations"]) == 12
    assert abs(res["Energy"] - g["E_converged"]) < 1e-8
    assert np.abs(res["Overlap"] - p["S"]).max() < 1e-9  # same molecule up to the rigid shift to the centre of mass


@pytest.fixture(scope="module")
def tetramer():
    """First four molecules of the w50 cluster, cc-pVDZ / cc-pVDZ-RIFIT (N = 100, Q = 384): the oracle's whole run on the
    CPU (own integrals, Schwarz rule, reference screened algorithm, SCF replay)."""
    import json
    import os
    from juliachem_jl_b200 import rhf
    geo = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "w50_geometry.json")))
    molecule = {"geometry": geo["geometry"][:36], "symbols": geo["symbols"][:12], "molecular_charge": 0}
    mol = rhf.run_basis(molecule, {"basis": "cc-pVDZ"})
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
    Bs = orc.calculate_B(np.tril(J), np.asfortranarray(T3.reshape(Q, N * N))[:, fn_mask.ravel(order="F")])
    nocc = mol["n_electrons"] // 2
    ref = sr.run_scf(S, H, E_nuc, nocc, lambda C, it: orc.df_rhf_fock_build([Bs], C, nocc, H, sd, "screened"))
    return mol, fn_mask, ref


@pytest.mark.parametrize("ndev", [1, 2])
def test_water_tetramer_from_w50_geometry(tetramer, ndev):
    """Everything on the GPU from the geometry on (one-electron integrals, metric, Schwarz mask, (P|mn), B, SCF loop)
    against the oracle doing the same run on the CPU: same mask, same iteration count, energy within 1e-8 Eh."""
    from juliachem_jl_b200 import rhf
    if jc.device_count() < ndev:
        pytest.skip("needs %d GPUs" % ndev)
    mol, fn_mask, ref = tetramer
    res = rhf.run_energy(mol, {"df_sigma": 1e-5}, num_devices=ndev)
    assert res["pairs_kept"] == int(fn_mask.sum()) and res["pairs_kept"] < fn_mask.size
    assert res["Converged?"] and ref["converged"]
    assert len(res["Iterations"]) == len(ref["iterations"])
    assert abs(res["Energy"] - ref["E"]) < 1e-8, (res["Energy"], ref["E"])
    assert np.abs(res["Density"] - ref["D"]).max() < 1e-6
