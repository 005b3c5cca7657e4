"""Generates the small data files the real-molecule runs use (run once, in the build container, where /root/reference
is mounted; the outputs are committed because the GPU box has no /root/reference):

  juliachem.jl_b200/data/basis_ccpvdz_OH.json   cc-pVDZ + cc-pVDZ-RIFIT shells of O and H exactly as the reference printed
                                                them (water_ccpvdz_out.log:43-170, already parsed into the golden fixture
                                                tests/golden/water_ccpvdz_rifit.json by make_water_golden.py).  The
                                                reference's own basis-set database records/bsed.h5 is a binary blob that
                                                is not part of the checkout.
  tests/golden/w50_geometry.json                symbols + Cartesian coordinates (Angstrom) of the 50-water cluster of
                                                example_inputs/w50-static.json (the input BASELINE.json's w50 config names).
"""
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"


def main():
    g = json.load(open(os.path.join(HERE, "water_ccpvdz_rifit.json")))
    lib = {"source": "water_ccpvdz_out.log:43-170 via tests/golden/water_ccpvdz_rifit.json", "basis": "cc-pVDZ",
           "auxiliary_basis": "cc-pVDZ-RIFIT", "atoms": {}}
    for atom_b, atom_a, z in zip(g["basis"], g["auxiliary_basis"], g["charges"]):
        sym = {8: "O", 1: "H"}[int(round(z))]
        if sym in lib["atoms"]:
            continue
        lib["atoms"][sym] = {"Z": int(round(z)),
                             "basis": [{"l": s["l"], "exps": s["exps"], "coefs": s["coefs"]} for s in atom_b["shells"]],
                             "auxiliary_basis": [{"l": s["l"], "exps": s["exps"], "coefs": s["coefs"]} for s in atom_a["shells"]]}
    out = os.path.join(ROOT, "juliachem.jl_b200", "data", "basis_ccpvdz_OH.json")
    json.dump(lib, open(out, "w"), indent=1)
    print("wrote", out)
    w = json.load(open(os.path.join(REF, "example_inputs", "w50-static.json")))
    geo = {"source": "example_inputs/w50-static.json (molecule.geometry, Angstrom, flat xyz per atom)",
           "symbols": w["molecule"]["symbols"], "geometry": w["molecule"]["geometry"],
           "molecular_charge": w["molecule"]["molecular_charge"]}
    out = os.path.join(HERE, "w50_geometry.json")
    json.dump(geo, open(out, "w"))
    print("wrote", out, len(geo["symbols"]), "atoms")


if __name__ == "__main__":
    main()
