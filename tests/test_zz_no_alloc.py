"""jcdf_fock_build_no_alloc (include/jcdf.h; no counterpart in the reference):
the dense Fock build with its scratch in the library's fixed __device__ array, called through the C-ABI on raw device
addresses and compared with the existing entry point — jcdf_create + jcdf_set_B + jcdf_fock_build on the same inputs — at the
1e-10 per element the other parity tests use, on calls in a row that share the array (a full-size call, then a smaller one
that must not see the first one's leftovers, then the first again).  Collected last on purpose."""
import os
import re

import numpy as np
import pytest

from helpers import ROOT, jc, oracle_fock, random_inputs, reference_layout_rows

gpu = pytest.mark.gpu
TOL = 1e-10
# (N, Q, occ): the array's maximum; water in 6-31G(2df,p) / cc-pVTZ-JKFIT (tests/golden/s10_631g2dfp_jkfit.json), ragged in
# every dimension; water in cc-pVDZ-like sizes with occ <= 4
CASES = [(48, 256, 16), (47, 166, 5), (25, 116, 3)]


def _existing_entry_point(B, C_occ, H):
    Q, N, _ = B.shape
    gd = jc.B200Data(N, Q, C_occ.shape[1], 1)
    gd.set_mask(jc.ScreeningData(None, N))
    gd.set_B(0, reference_layout_rows(B, range(Q), None))
    F = gd.fock_build(C_occ, H)[0].copy()
    gd.close()
    return F


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=np.float64).ravel(order="F"))).cuda()


def _no_alloc(B, C_occ, H, pad=0):
    """One call through the C-ABI; B, C_occ (leading dimensions padded by `pad`) and H are copied to the device by torch —
    the caller's allocations — and the library allocates nothing."""
    import torch
    Q, N, _ = B.shape
    occ = C_occ.shape[1]
    B2 = np.zeros((Q + pad, N * N), order="F")
    B2[:Q] = B.reshape((Q, N * N), order="F")
    Cp = np.zeros((N + pad, occ), order="F")
    Cp[:N] = C_occ
    dB, dC, dH = _dev(B2), _dev(Cp), (None if H is None else _dev(H))
    dF = torch.full((N * N,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    jc.fock_build_no_alloc(N, Q, occ, dB.data_ptr(), Q + pad, dC.data_ptr(), N + pad, None if dH is None else dH.data_ptr(),
                           dF.data_ptr())
    torch.cuda.synchronize()
    return dF.cpu().numpy().reshape((N, N), order="F")


@gpu
def test_no_alloc_matches_existing_entry_point_on_calls_in_a_row():
    import torch
    torch.cuda.set_device(0)
    inputs = [random_inputs(N, Q, occ, seed=900 + k) for k, (N, Q, occ) in enumerate(CASES)]
    expected = [_existing_entry_point(*x) for x in inputs]
    first = None
    # full size, then smaller shapes on the same (uncleared) array, then the full size again
    for k in (0, 1, 2, 0, 1):
        B, C_occ, H = inputs[k]
        F = _no_alloc(B, C_occ, H, pad=3 * (k % 2))
        assert np.isfinite(F).all() and np.array_equal(F, F.T)
        err = float(np.abs(F - expected[k]).max())
        print("no_alloc N, Q, occ = %s: max|F - F_existing| = %.3e on max|F| = %.3f" % (CASES[k], err, np.abs(F).max()))
        assert err < TOL
        assert np.abs(F - oracle_fock(B, C_occ, H, None)).max() < TOL
        if k == 0:
            if first is None:
                first = F
            else:
                assert np.array_equal(F, first)  # fixed summation order: the repeat is bit-identical
    # H = NULL: the two-electron part alone, as the existing entry point returns it for H = None
    B, C_occ, H = inputs[1]
    F2 = _no_alloc(B, C_occ, None)
    assert np.abs(F2 - (expected[1] - H)).max() < TOL


@gpu
def test_no_alloc_two_calls_back_to_back_on_one_stream():
    """Two calls enqueued without a synchronisation in between share the array in stream order."""
    import torch
    torch.cuda.set_device(0)
    (Ba, Ca, Ha), (Bb, Cb, Hb) = random_inputs(40, 200, 9, seed=77), random_inputs(33, 70, 16, seed=78)
    exp_a, exp_b = _existing_entry_point(Ba, Ca, Ha), _existing_entry_point(Bb, Cb, Hb)
    st = torch.cuda.Stream()
    da = [_dev(Ba.reshape((200, 1600), order="F")), _dev(Ca), _dev(Ha), torch.empty(1600, dtype=torch.float64, device="cuda")]
    db = [_dev(Bb.reshape((70, 1089), order="F")), _dev(Cb), _dev(Hb), torch.empty(1089, dtype=torch.float64, device="cuda")]
    torch.cuda.synchronize()
    jc.fock_build_no_alloc(40, 200, 9, da[0].data_ptr(), 200, da[1].data_ptr(), 40, da[2].data_ptr(), da[3].data_ptr(),
                           st.cuda_stream)
    jc.fock_build_no_alloc(33, 70, 16, db[0].data_ptr(), 70, db[1].data_ptr(), 33, db[2].data_ptr(), db[3].data_ptr(),
                           st.cuda_stream)
    st.synchronize()
    Fa = da[3].cpu().numpy().reshape((40, 40), order="F")
    Fb = db[3].cpu().numpy().reshape((33, 33), order="F")
    assert np.abs(Fa - exp_a).max() < TOL and np.abs(Fb - exp_b).max() < TOL


@gpu
@pytest.mark.parametrize("N,Q,occ", [(49, 10, 2), (10, 257, 2), (20, 10, 17), (0, 10, 1), (4, 10, 5)])
def test_no_alloc_refuses_sizes_beyond_the_fixed_array(N, Q, occ):
    import torch
    d = torch.zeros(8, dtype=torch.float64, device="cuda")
    with pytest.raises(jc.JCDFError) as e:
        jc.fock_build_no_alloc(N, Q, occ, d.data_ptr(), max(Q, 1), d.data_ptr(), max(N, 1), None, d.data_ptr())
    assert "fixed scratch array" in str(e.value) or "occ > N" in str(e.value)


def test_no_alloc_source_calls_no_allocator_and_takes_the_symbol_address():
    """CPU check of the translation unit itself: no allocating call on its path, the array is a file-scope static __device__
    __align__(256) array whose DEVICE address comes from a checked cudaGetSymbolAddress, and its byte size is the documented
    one."""
    src = open(os.path.join(ROOT, "juliachem.jl_b200", "csrc", "no_alloc.cu")).read()
    code = re.sub(r"//[^\n]*", "", src)
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMemPool", "cublas", "cusolver", "torch", "new ", "std::vector"):
        assert banned not in code, banned
    assert re.search(r"static __device__ __align__\(256\) double g_no_alloc_scratch\[", code)
    assert "JC_CUDA(cudaGetSymbolAddress((void**)&p, g_no_alloc_scratch))" in code
    # kernels receive the looked-up pointer, never the symbol: it appears in its definition and in that one call only
    assert code.count("g_no_alloc_scratch") == 2
    hdr = open(os.path.join(ROOT, "include", "jcdf.h")).read()
    n, q, o = (int(re.search(r"#define JCDF_NO_ALLOC_MAX_%s (\d+)" % k, hdr).group(1)) for k in ("N", "Q", "OCC"))
    assert (n, q, o) == (48, 256, 16)
    assert 8 * (o * n * q + q + (q // 32) * n * n) == 1722368 and "1 722 368 bytes" in hdr
