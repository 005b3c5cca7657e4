"""Third pin of the oracle, against the values the reference's OWN test suite asserts (test/runtests.jl:62-63): the S22 water
dimer in 6-31G(2df,p) / cc-pVTZ-JKFIT.  The reference requires its RHF energy to equal the GAMESS value of
test/s22_gamess_values.json and its DF-RHF energy to lie within 1.5e-3 Eh of it; the oracle's DF-RHF energy — own integrals,
real Schwarz mask (78 % of the pairs survive), the default screened algorithm — must meet the same criterion, and in fact
differs from the independent GAMESS number by 2.7e-5 Eh, the density-fitting error of cc-pVTZ-JKFIT.  Fixture
tests/golden/s22_hco.json (tests/golden/make_s22_golden.py).  The other nine H/C/O complexes run on the GPU
(tests/test_gpu_s22_reference_suite.py).  CPU only."""
import json
import os

from oracle import df_oracle as orc
from oracle import molecule_case as mc
from oracle import scf_replay as sr

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = json.load(open(os.path.join(HERE, "golden", "s22_hco.json")))


def oracle_case(m, basis_name):
    """Golden-style `case` dict of a JSON-contract molecule, through the host mirror of JCBasis.run (bohr, centre of mass)."""
    from juliachem_jl_b200 import rhf
    mol = rhf.run_basis(m, {"basis": basis_name})
    lib = rhf.load_basis_library(basis_name)
    return {"geometry": [{"symbol": s, "xyz_bohr": list(map(float, c))} for s, c in zip(mol["symbols"], mol["centers"])],
            "charges": [int(z) for z in mol["charges"]], "n_electrons": mol["n_electrons"],
            "basis": [{"symbol": s, "shells": lib[s]["basis"]} for s in mol["symbols"]],
            "auxiliary_basis": [{"symbol": s, "shells": lib[s]["auxiliary_basis"]} for s in mol["symbols"]]}


def test_fixture_is_the_references_suite():
    assert GOLDEN["df_vs_rhf_atol"] == 0.0015 and len(GOLDEN["molecules"]) == 10
    assert GOLDEN["molecules"]["11"]["E_RHF_GAMESS"] == -461.4497600413 and len(GOLDEN["molecules"]["11"]["symbols"]) == 24


def test_water_dimer_df_energy_meets_the_references_criterion():
    m = GOLDEN["molecules"]["2"]
    p = mc.build_problem(oracle_case(m, GOLDEN["basis"]), 1e-5, 1, min(8, os.cpu_count() or 1))
    assert (p["N"], p["Q"], p["nocc"]) == (94, 332, 10) and 0.7 < p["fn_mask"].mean() < 0.85
    H, nocc, sd = p["H"], p["nocc"], p["sd"]
    res = sr.run_scf(p["S"], H, p["E_nuc"], nocc, lambda C, it: orc.df_rhf_fock_build(p["B_shards"], C, nocc, H, sd, "screened"),
                     max_iterations=100)
    assert res["converged"]
    assert abs(res["E"] - m["E_RHF_GAMESS"]) <= GOLDEN["df_vs_rhf_atol"]   # runtests.jl:63
    assert 0 < res["E"] - m["E_RHF_GAMESS"] < 1e-4                          # DF error of cc-pVTZ-JKFIT: 2.7e-5 Eh
