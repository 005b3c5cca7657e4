#!/usr/bin/env python
"""The reference's own test suite on the B200 path (test/runtests.jl:62-63): every S22 complex of H, C, O in 6-31G(2df,p) /
cc-pVTZ-JKFIT as a whole DF-RHF run on the device, energy against the GAMESS RHF value the reference tests against, with the
reference's tolerance for DF vs RHF (1.5e-3 Eh).  Prints / writes one record per complex.
    python tools/s22_suite.py [num_devices=1] [out.json]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(num_devices=1, only=None):
    from juliachem_jl_b200 import rhf
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "s22_hco.json")))
    out = []
    for key in sorted(g["molecules"], key=int):
        if only and int(key) not in only:
            continue
        m = g["molecules"][key]
        mol = rhf.run_basis(m, {"basis": g["basis"]})
        t0 = time.perf_counter()
        res = rhf.run_energy(mol, {"niter": 100}, num_devices=num_devices)
        wall = time.perf_counter() - t0
        fock = sorted(v for k, v in res["Timings"]["jc"].items() if k.startswith("fock_time"))
        out.append({"s22": int(key), "name": m["name"], "N": res["N"], "Naux": res["Q"], "occ": res["occ"],
                    "mask_survival": res["pairs_kept"] / float(res["N"]) ** 2, "E_df_b200": res["Energy"],
                    "E_rhf_gamess": m["E_RHF_GAMESS"], "df_minus_rhf": res["Energy"] - m["E_RHF_GAMESS"],
                    "converged": bool(res["Converged?"]), "iterations": len(res["Iterations"]),
                    "fock_ms_median": round(fock[len(fock) // 2] * 1e3, 3), "wall_seconds": round(wall, 2)})
        print(json.dumps(out[-1]), flush=True)
    return {"criterion": "test/runtests.jl:63  |E_RHF - E_DF| <= 1.5e-3 Eh (E_RHF ~ GAMESS, :62)", "atol": g["df_vs_rhf_atol"],
            "devices": num_devices, "results": out}


if __name__ == "__main__":
    nd = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    rec = run(nd)
    path = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "r02_s22_suite.json")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    json.dump(rec, open(path, "w"), indent=1)
    bad = [r for r in rec["results"] if not (r["converged"] and abs(r["df_minus_rhf"]) <= rec["atol"])]
    sys.exit(1 if bad else 0)
