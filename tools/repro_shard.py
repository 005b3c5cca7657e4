"""Emulates rank r of a W-rank run on one GPU: real w50, shard [Q r / W, Q (r+1) / W), one Fock build; reports whether
the partial Fock matrix is finite and symmetric and its checksum.  python tools/repro_shard.py W [dense]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import juliachem_jl_b200 as jc
from juliachem_jl_b200 import workloads as wl, rhf, multi

W = int(sys.argv[1])
dense = len(sys.argv) > 2
w = wl.WORKLOADS["w50_dense" if dense else "w50"]
C, H = wl.occupied_coefficients(w), wl.core_hamiltonian(w)
total = np.zeros((w.N, w.N))
for r in range(W):
    gd = jc.B200Data(w.N, w.Q, w.occ, 1)
    gd.set_shards(multi.rank_shards(w.Q, W, r, 1))
    if dense:
        gd.set_mask(jc.ScreeningData(None, w.N))
        gd.fill_synthetic_B(w.seed, w.decay_length)
    else:
        mol = rhf.run_basis(wl.real_molecule(w), {"basis": "cc-pVDZ"})
        gd.set_basis(0, *mol["primary"].args()); gd.set_basis(1, *mol["auxiliary"].args())
        gd.compute_two_center(False); gd.compute_schwarz_mask(1e-5); gd.compute_B()
    F, st = gd.fock_build(C, H if r == 0 else None)
    if r > 0:
        pass
    print("rank", r, "finite", bool(np.isfinite(F).all()), "sym", bool(np.array_equal(F, F.T)), "sum", float(np.nansum(F)),
          "nan count", int(np.isnan(F).sum()), {k: round(v * 1e3, 2) for k, v in st.items() if k in "WKJV"}, flush=True)
    total += np.nan_to_num(F)
    gd.close()
print("total checksum", float(total.sum()), float(np.abs(total).max()))
