#!/usr/bin/env python
"""Third reference-held golden: the reference's OWN test suite.  test/runtests.jl:1-97 runs every S22 complex
(example_inputs/S22/NN_MP2.json, basis 6-31G(2df,p)) as conventional RHF and as DF-RHF with cc-pVTZ-JKFIT and asserts
    RHF energy  ≈ GAMESS value (test/s22_gamess_values.json)                    (:62)
    |RHF energy - DF-RHF energy| <= 1.5e-3 Eh                                   (:63)
This script collects, for the S22 complexes made of H, C and O only (the elements whose 6-31G(2df,p) / cc-pVTZ-JKFIT data the
snapshot's logs print, tests/golden/make_s10_golden.py), the input geometry (Angstrom, as in the JSON contract) and the GAMESS
energy into tests/golden/s22_hco.json.  Run in the build container only:   python tests/golden/make_s22_golden.py"""
import json
import os

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
NAMES = {2: "water dimer", 3: "formic acid dimer", 8: "methane dimer", 9: "ethene dimer", 10: "benzene - methane",
         11: "benzene dimer (parallel displaced)", 16: "ethene - ethyne", 17: "benzene - water", 20: "benzene dimer (T-shaped)",
         22: "phenol dimer"}


def main():
    gamess = json.load(open(os.path.join(REF, "test", "s22_gamess_values.json")))
    out = {"source": "example_inputs/S22/NN_MP2.json + test/s22_gamess_values.json (j-hayes/JuliaChem.jl @ 2064729); criterion "
                     "test/runtests.jl:62-63", "basis": "6-31G(2df,p)", "auxiliary_basis": "cc-pVTZ-JKFIT",
           "df_vs_rhf_atol": 0.0015, "molecules": {}}
    for i in range(1, 23):
        inp = json.load(open(os.path.join(REF, "example_inputs", "S22", "%02d_MP2.json" % i)))
        mol = inp["molecule"]
        if not set(mol["symbols"]) <= {"H", "C", "O"}:
            continue
        assert inp["model"]["basis"] == "6-31G(2df,p)" and i in NAMES
        out["molecules"][str(i)] = {"name": NAMES[i], "symbols": mol["symbols"], "geometry": mol["geometry"],
                                    "molecular_charge": mol["molecular_charge"], "E_RHF_GAMESS": gamess[str(i)]["Energy"],
                                    "homo_lumo_gap_GAMESS": gamess[str(i)]["HOMO-LUMO Gap"]}
    assert sorted(map(int, out["molecules"])) == sorted(NAMES)
    p = os.path.join(HERE, "s22_hco.json")
    json.dump(out, open(p, "w"))
    print("wrote", p, os.path.getsize(p), "bytes,", len(out["molecules"]), "complexes")


if __name__ == "__main__":
    main()
