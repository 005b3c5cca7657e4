#!/usr/bin/env python
"""Real-molecule run of the B200 path: the first n molecules of the 50-water cluster (tests/golden/w50_geometry.json) in
cc-pVDZ / cc-pVDZ-RIFIT, everything on the GPU from the geometry on.  Prints one JSON line with the energy, iteration
table and the wall time of every stage.   python tools/run_w50_real.py [n_molecules=50] [num_devices=1] [df_sigma=1e-5]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from juliachem_jl_b200 import rhf  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 50
    ndev = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    sigma = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-5
    geo = json.load(open(os.path.join(ROOT, "tests", "golden", "w50_geometry.json")))
    mol = rhf.run_basis({"geometry": geo["geometry"][:9 * n], "symbols": geo["symbols"][:3 * n], "molecular_charge": 0},
                        {"basis": "cc-pVDZ"})
    mode = sys.argv[4] if len(sys.argv) > 4 else "exchange"
    res = rhf.run_energy(mol, {"df_sigma": sigma, "b_formation": mode}, num_devices=ndev)
    li = res["layout"]
    t = res["Timings"]
    fock = [v for k, v in t["jc"].items() if k.startswith("fock_time")]
    print(json.dumps({"molecules": n, "N": res["N"], "Q": res["Q"], "occ": res["occ"], "devices": ndev, "df_sigma": sigma,
                      "pairs_kept": res["pairs_kept"], "mask_survival": res["pairs_kept"] / float(res["N"]) ** 2,
                      "tiles_kept": li.tiles_kept, "tiles_total": li.tiles_total, "E": res["Energy"],
                      "converged": res["Converged?"], "iterations": res["Iterations"],
                      "seconds": {k: round(v, 4) for k, v in t.items() if k != "jc"},
                      "fock_ms_median": round(sorted(fock)[len(fock) // 2] * 1e3, 3)}))


if __name__ == "__main__":
    main()
