"""The reference's own acceptance test on the B200 path.  test/runtests.jl:1-97 runs every S22 complex in 6-31G(2df,p) as
conventional RHF (must equal the GAMESS energy of test/s22_gamess_values.json, :62) and as DF-RHF with cc-pVTZ-JKFIT (must be
within 1.5e-3 Eh of the RHF energy, :63).  Here: the ten complexes made of H, C and O (the elements whose basis data the
snapshot holds; fixture tests/golden/s22_hco.json), each a whole DF-RHF run on the device from the geometry on — device
integrals incl. primary f and auxiliary g shells, device Schwarz masks of real molecules, the screened Fock build, DIIS — with the
reference's tolerance against the GAMESS value.  The water dimer additionally against the oracle's SCF to 1e-8 Eh."""
import json
import os
import sys

import numpy as np
import pytest

from helpers import ROOT, jc, orc

pytestmark = pytest.mark.gpu
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "s22_hco.json")))


@pytest.mark.parametrize("key", sorted(GOLDEN["molecules"], key=int))
def test_s22_df_energy_within_the_references_tolerance_of_gamess(key):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import s22_suite
    r = s22_suite.run(1, only=[int(key)])["results"][0]
    assert r["converged"], r
    assert abs(r["df_minus_rhf"]) <= GOLDEN["df_vs_rhf_atol"], r   # runtests.jl:63
    assert abs(r["df_minus_rhf"]) <= 3e-4, r   # what cc-pVTZ-JKFIT actually delivers here: 8e-6 .. 2.2e-4 (profiles/r02_s22_suite.json)


def test_s22_water_dimer_matches_the_oracle_scf():
    from juliachem_jl_b200 import rhf
    from oracle import molecule_case as mc
    from oracle import scf_replay as sr
    from test_oracle_s22_reference_suite import oracle_case
    m = GOLDEN["molecules"]["2"]
    mol = rhf.run_basis(m, {"basis": GOLDEN["basis"]})
    case = oracle_case(m, GOLDEN["basis"])
    p = mc.build_problem(case, 1e-5, 1, min(8, os.cpu_count() or 1))
    H, nocc, sd = p["H"], p["nocc"], p["sd"]
    ref = sr.run_scf(p["S"], H, p["E_nuc"], nocc, lambda C, it: orc.df_rhf_fock_build(p["B_shards"], C, nocc, H, sd, "screened"),
                     max_iterations=100)
    res = rhf.run_energy(mol, {"niter": 100})
    assert res["Converged?"] and ref["converged"] and len(res["Iterations"]) == len(ref["iterations"])
    assert abs(res["Energy"] - ref["E"]) <= 1e-8
    assert abs(ref["E"] - m["E_RHF_GAMESS"]) <= GOLDEN["df_vs_rhf_atol"]
