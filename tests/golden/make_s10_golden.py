#!/usr/bin/env python
"""Second reference-held golden: /root/reference/test/s10_new_algo-3-20.log — the reference's own density-fitting test
script (test/density-fitting-test.jl, run by test/density-fitting-test-3.sh on 3 MPI ranks x 20 threads) in
6-31G(2df,p) / cc-pVTZ-JKFIT with the DEFAULT (screened) contraction mode and static Q sharding.  What the log holds:

  run 1 (:26-281)   water, "Density Fitting: true, Contraction Mode: default", 3 ranks: 13 printed iteration lines
                    (:264-276) and "Total SCF Energy: -75.9905373885 h" (:281).  N = 47, Naux = 166 (auxiliary shells up
                    to g on O), 10 electrons.
  run 2 (:337-479)  the same water by conventional four-index RHF ("Density Fitting: false"): 14 iteration lines and
                    E = -75.9905503208 (:479) — what test/runtests.jl:63 compares DF against (1.5e-3 Eh).
  run 3 (:624-1633) S22 system 10 (benzene-methane, C7H10; geometry :1580-1596) DF, same modes, 3 ranks: N = 297,
                    Naux = 1022, 52 electrons; the job was terminated after two printed iterations (:1632-1633) — their
                    energies still pin the first two screened + sharded Fock builds on a carbon system.

The basis printout carries 6 decimals, so energies are reproducible to ~1e-5 Eh, not 1e-8 (stated in the tests).
Outputs (committed; the GPU box has no /root/reference):
  tests/golden/s10_631g2dfp_jkfit.json                     both DF cases + the RHF energy
  juliachem.jl_b200/data/basis_631g2dfp_jkfit_HCO.json     6-31G(2df,p) + cc-pVTZ-JKFIT for H, C, O as printed
Run in the build container only:   python tests/golden/make_s10_golden.py
"""
import json
import os
import re

LOG = "/root/reference/test/s10_new_algo-3-20.log"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
AM = {"S": 0, "P": 1, "D": 2, "F": 3, "G": 4}
ZNUC = {"H": 1, "C": 6, "N": 7, "O": 8}


def parse_basis(lines):
    """Atoms of one 'Printing ... basis set' block.  L shells are printed as 'L (s)' then 'L (p)' rows and become two
    shells, s first (JCBasis.jl:245-290)."""
    atoms, cur_atom, cur = [], None, None
    for ln in lines:
        m = re.match(r"Atom #(\d+) \((\w+)\):", ln.strip())
        if m:
            cur_atom = {"symbol": m.group(2), "shells": []}
            atoms.append(cur_atom)
            cur = None
            continue
        f = ln.split()
        if len(f) == 5 and f[1] in AM and f[0].isdigit():
            key, l = (int(f[0]), f[1]), AM[f[1]]
        elif len(f) == 6 and f[1] == "L" and f[2] in ("(s)", "(p)") and f[0].isdigit():
            key, l = (int(f[0]), f[2]), 0 if f[2] == "(s)" else 1
            f = [f[0], f[1]] + f[3:]
        else:
            continue
        if cur is None or cur["key"] != key:
            cur = {"key": key, "l": l, "exps": [], "coefs": []}
            cur_atom["shells"].append(cur)
        cur["exps"].append(float(f[3]))
        cur["coefs"].append(float(f[4]))
    for a in atoms:
        for s in a["shells"]:
            del s["key"]
    return atoms


def section(lines, start):
    """One run of the script: (basis, aux or None, geometry in bohr, options, iteration rows, E_total or None)."""
    i = start
    while "Printing basis set..." not in lines[i]:
        i += 1
    j = i
    while "Printing basis set metadata" not in lines[j]:
        j += 1
    blk = lines[i:j]
    k = next((n for n, ln in enumerate(blk) if "Printing Auxillary basis set" in ln), None)
    basis = parse_basis(blk[:k] if k is not None else blk)
    aux = parse_basis(blk[k:]) if k is not None else None
    while "in xyz format" not in lines[j]:
        j += 1
    j += 1
    while not lines[j].strip().isdigit():
        j += 1
    natoms = int(lines[j].strip())
    geom = []
    j += 1
    while len(geom) < natoms:
        f = lines[j].split()
        if len(f) == 4 and f[0] in ZNUC:
            geom.append({"symbol": f[0], "xyz_bohr": [float(x) for x in f[1:4]]})
        j += 1
    opts = {}
    while not lines[j].startswith("Iter        Energy"):
        m = re.match(r"([A-Za-z ]+): (.+)", lines[j])
        if m:
            opts[m.group(1).strip()] = m.group(2).strip()
        j += 1
    iters, e_tot = [], None
    j += 1
    while j < len(lines) and "END RESTRICTED CLOSED-SHELL" not in lines[j]:
        f = lines[j].split()
        if len(f) == 4 and f[0].isdigit() and "." in f[1]:
            iters.append({"iter": int(f[0]), "E": float(f[1]), "dE": float(f[2]), "D_rms": float(f[3])})
        m = re.match(r"Total SCF Energy: (-?[\d.]+) h", lines[j])
        if m:
            e_tot = float(m.group(1))
        j += 1
    return basis, aux, geom, opts, iters, e_tot, j


def case(basis, aux, geom, opts, iters, e_tot, lines_cited):
    nel = sum(ZNUC[g["symbol"]] for g in geom)
    return {"log_lines": lines_cited, "options": opts, "geometry": geom, "charges": [ZNUC[g["symbol"]] for g in geom],
            "n_electrons": nel, "basis": basis, "auxiliary_basis": aux, "iterations": iters, "E_converged": e_tot}


def main():
    lines = open(LOG).read().splitlines()
    b1, a1, g1, o1, it1, e1, end1 = section(lines, 0)
    b2, a2, g2, o2, it2, e2, end2 = section(lines, end1)
    b3, a3, g3, o3, it3, e3, _ = section(lines, end2)
    assert o1["Density Fitting"] == "true" and o1["Contraction Mode"] == "default" and len(it1) == 13 and e1 == -75.9905373885
    assert o2["Density Fitting"] == "false" and len(it2) == 14 and e2 == -75.9905503208
    assert o3["Density Fitting"] == "true" and len(g3) == 17 and len(it3) == 2 and e3 is None
    assert g1 == g2 and b1 == b2

    def nfun(atoms):
        return sum((s["l"] + 1) * (s["l"] + 2) // 2 for a in atoms for s in a["shells"])
    assert (nfun(b1), nfun(a1), nfun(b3), nfun(a3)) == (47, 166, 297, 1022)
    out = {
        "source": "test/s10_new_algo-3-20.log (j-hayes/JuliaChem.jl @ 2064729); script test/density-fitting-test.jl, "
                  "launcher test/density-fitting-test-3.sh: mpiexecjl -np 3, 20 threads per rank",
        "basis_name": "6-31G(2df,p)", "auxiliary_basis_name": "cc-pVTZ-JKFIT", "n_ranks": 3,
        "settings": {"contraction_mode": "default", "load": "static", "guess": "hcore", "dele": 1e-6, "rmsd": 1e-6,
                     "ndiis": 10, "df_sigma": 1e-5},
        "water": case(b1, a1, g1, o1, it1, e1, "26-281"),
        "water_rhf_conventional": {"log_lines": "337-479", "iterations": it2, "E_converged": e2},
        "benzene_methane": case(b3, a3, g3, o3, it3, e3, "624-1633"),
    }
    p = os.path.join(HERE, "s10_631g2dfp_jkfit.json")
    json.dump(out, open(p, "w"))
    print("wrote", p, os.path.getsize(p), "bytes")
    # product data file: one entry per element, shells exactly as printed
    lib = {"source": "test/s10_new_algo-3-20.log:43-205 (O, H) and :641-1560 (C) via tests/golden/make_s10_golden.py",
           "basis": "6-31G(2df,p)", "auxiliary_basis": "cc-pVTZ-JKFIT", "atoms": {}}
    for atoms_b, atoms_a in ((b1, a1), (b3, a3)):
        for ab, aa in zip(atoms_b, atoms_a):
            assert ab["symbol"] == aa["symbol"]
            entry = {"Z": ZNUC[ab["symbol"]], "basis": ab["shells"], "auxiliary_basis": aa["shells"]}
            if ab["symbol"] in lib["atoms"]:
                assert lib["atoms"][ab["symbol"]] == entry, "the log prints one element two different ways"
            lib["atoms"][ab["symbol"]] = entry
    assert sorted(lib["atoms"]) == ["C", "H", "O"]
    p = os.path.join(ROOT, "juliachem.jl_b200", "data", "basis_631g2dfp_jkfit_HCO.json")
    json.dump(lib, open(p, "w"), indent=1)
    print("wrote", p)


if __name__ == "__main__":
    main()
