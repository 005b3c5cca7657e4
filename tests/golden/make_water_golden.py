#!/usr/bin/env python
"""Extracts the golden fixture tests/golden/water_ccpvdz_rifit.json from the reference's own logged run
/root/reference/water_ccpvdz_out.log (water, cc-pVDZ / cc-pVDZ-RIFIT, dense CPU DF-RHF, hcore guess):

  * the basis and auxiliary basis exactly as printed (:43-170), the geometry in bohr (:195-197);
  * the per-iteration table (iter, E, dE, D_rms) (:253-523) and the converged energy (:563);
  * the unlabeled 25 x 25 matrix printed before every iteration line (8 significant decimals).

Run in the build container only (it reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_water_golden.py
"""
import json
import os
import re

LOG = "/root/reference/water_ccpvdz_out.log"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "water_ccpvdz_rifit.json")
AM = {"S": 0, "P": 1, "D": 2, "F": 3, "G": 4}
ZNUC = {"H": 1, "O": 8, "C": 6, "N": 7}


def parse_basis(lines, start_marker, end_markers):
    atoms, cur_atom, cur_shell = [], None, None
    active = False
    for ln in lines:
        if start_marker in ln:
            active = True
            continue
        if active and any(m in ln for m in end_markers):
            break
        if not active:
            continue
        m = re.match(r"Atom #(\d+) \((\w+)\):", ln.strip())
        if m:
            cur_atom = {"symbol": m.group(2), "shells": []}
            atoms.append(cur_atom)
            cur_shell = None
            continue
        f = ln.split()
        if len(f) == 5 and f[1] in AM and f[0].isdigit():
            sid = int(f[0])
            if cur_shell is None or cur_shell["id"] != sid:
                cur_shell = {"id": sid, "l": AM[f[1]], "exps": [], "coefs": []}
                cur_atom["shells"].append(cur_shell)
            cur_shell["exps"].append(float(f[3]))
            cur_shell["coefs"].append(float(f[4]))
    for a in atoms:
        for s in a["shells"]:
            del s["id"]
    return atoms


def main():
    lines = open(LOG).read().splitlines()
    basis = parse_basis(lines, "Printing basis set...", ["Printing Auxillary basis set"])
    aux = parse_basis(lines, "Printing Auxillary basis set...", ["Printing basis set metadata"])
    geom = []
    for i, ln in enumerate(lines):
        if ln.strip() == "3" and "xyz" in lines[i - 3]:
            for g in lines[i + 2:i + 5]:
                f = g.split()
                geom.append({"symbol": f[0], "xyz_bohr": [float(x) for x in f[1:4]]})
    iters, mats, cur = [], [], []
    in_scf = False
    for ln in lines:
        if ln.startswith("Iter        Energy"):
            in_scf = True
            continue
        if not in_scf:
            continue
        if "SCF calculation has converged" in ln:
            break
        f = ln.split()
        if len(f) == 25:
            cur.append([float(x) for x in f])
        elif len(f) == 5 and f[0].isdigit():
            iters.append({"iter": int(f[0]), "E": float(f[1]), "dE": float(f[2]), "D_rms": float(f[3])})
            mats.append(cur[-25:])
            cur = []
    e_final = None
    for ln in lines:
        m = re.match(r"Total SCF Energy: (-?[\d.]+) h", ln)
        if m:
            e_final = float(m.group(1))
    assert len(basis) == 3 and len(aux) == 3 and len(geom) == 3 and len(iters) == 11 and all(len(m) == 25 for m in mats)
    out = {
        "source": "water_ccpvdz_out.log (j-hayes/JuliaChem.jl @ 2064729)",
        "settings": {"contraction_mode": "dense", "guess": "hcore", "dele": 1e-6, "rmsd": 1e-6, "ndiis": 10,
                     "n_electrons": 10, "N": 25, "Naux": 96},
        "geometry": geom, "charges": [ZNUC[g["symbol"]] for g in geom],
        "basis": basis, "auxiliary_basis": aux,
        "iterations": iters, "E_converged": e_final,
        "matrix_iter1": mats[0], "matrix_last": mats[-1],
    }
    json.dump(out, open(OUT, "w"))
    print("wrote", OUT, os.path.getsize(OUT), "bytes;", len(iters), "iterations, E =", e_final)


if __name__ == "__main__":
    main()
