import json, os, sys
import numpy as np
sys.path.insert(0, os.getcwd())
from juliachem_jl_b200 import rhf
from juliachem_jl_b200.df import B200Data
geo = json.load(open("tests/golden/w50_geometry.json"))
for n in (20, 50):
    mol = rhf.run_basis({"geometry": geo["geometry"][:9*n], "symbols": geo["symbols"][:3*n], "molecular_charge": 0}, {"basis": "cc-pVDZ"})
    prim, aux = mol["primary"], mol["auxiliary"]
    gd = B200Data(prim.n_functions, aux.n_functions, mol["n_electrons"]//2, 1)
    gd.set_basis(0, *prim.args()); gd.set_basis(1, *aux.args())
    gd.compute_two_center(False)
    mask, P = gd.compute_schwarz_mask(1e-5)
    np.save("gpurun_out/mask_w%d.npy" % n, np.packbits(mask, axis=None))
    print(n, P, mask.shape)
    gd.close()
