"""dca_adv_moments_static (include/dca_b200.h): the allocation-free twin of dca_adv_moments, its per-CTA partials in a
library-owned `static __device__` array.  No counterpart in the reference.
Collected last on purpose: a failure here hides nothing behind it."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _moments(lib, torch, ret, val, valid, vgroup, static):
    stats = torch.full((3,), -7.0, dtype=torch.float64, device=ret.device)  # poisoned: the kernel must write all three
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    vp = C.c_void_p(valid.data_ptr()) if valid is not None else None
    n = ret.numel()
    if static:
        rc = lib.dca_adv_moments_static(C.c_void_p(ret.data_ptr()), C.c_void_p(val.data_ptr()), vp, vgroup, n,
                                        C.c_void_p(stats.data_ptr()), stream)
    else:
        scratch = torch.empty(1776, dtype=torch.float64, device=ret.device)
        rc = lib.dca_adv_moments(C.c_void_p(ret.data_ptr()), C.c_void_p(val.data_ptr()), vp, vgroup, n,
                                 C.c_void_p(scratch.data_ptr()), C.c_void_p(stats.data_ptr()), stream)
    assert rc == 0, lib.dca_last_error(None)
    torch.cuda.synchronize()
    return stats.cpu().numpy()


def test_adv_moments_static_matches_adv_moments_twice_in_a_row():
    import torch
    from hmp2g_b200 import _lib
    lib = _lib.load()
    rng = np.random.RandomState(11)
    # two calls in a row on DIFFERENT batches (the second smaller and masked): whatever the shared array kept from
    # the first call must not leak into the second.  Same kernels, same fixed summation order -> bit-equal.
    cases = [((180, 64, 50), None), ((7, 3, 5), 5), ((180, 64, 50), 50), ((1, 1, 2), None)]
    for shape, vgroup in cases:
        ret = torch.from_numpy((rng.randn(*shape) * 3 + 1).astype(np.float32)).cuda()
        val = torch.from_numpy(rng.randn(*shape).astype(np.float32)).cuda()
        valid = None
        if vgroup is not None:
            rows = rng.rand(*shape[:2]) < 0.6
            rows[0, 0] = True
            valid = torch.from_numpy(rows.astype(np.uint8)).cuda()
        want = _moments(lib, torch, ret, val, valid, vgroup or 1, static=False)
        got = _moments(lib, torch, ret, val, valid, vgroup or 1, static=True)
        assert np.array_equal(got, want), (shape, got, want)
        again = _moments(lib, torch, ret, val, valid, vgroup or 1, static=True)
        assert np.array_equal(again, want), (shape, again, want)
        # and against the plain formula (1e-5, the tolerance of test_cuda_advantage_normalisation_matches_numpy_formula)
        adv = (ret - val).double()
        if valid is not None:
            adv = adv[valid.bool()]
        ref = np.array([adv.sum().item(), (adv * adv).sum().item(), float(adv.numel())])
        assert np.allclose(got, ref, rtol=1e-5, atol=1e-5), (shape, got, ref)


def test_adv_moments_static_allocates_nothing_and_rejects_oversize():
    import torch
    from hmp2g_b200 import _lib
    lib = _lib.load()
    ret = torch.randn(4096, device='cuda')
    val = torch.randn(4096, device='cuda')
    stats = torch.zeros(3, dtype=torch.float64, device='cuda')
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    args = (C.c_void_p(ret.data_ptr()), C.c_void_p(val.data_ptr()), None, 1)
    assert lib.dca_adv_moments_static(*args, 4096, C.c_void_p(stats.data_ptr()), stream) == 0  # address resolved
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(8):
        assert lib.dca_adv_moments_static(*args, 4096, C.c_void_p(stats.data_ptr()), stream) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0  # no device allocation behind the call
    too_many = (1 << 28) + 1  # DCA_ADV_STATIC_MAX_N + 1: refused before anything is launched or read
    rc = lib.dca_adv_moments_static(*args, too_many, C.c_void_p(stats.data_ptr()), stream)
    assert rc == -1 and b'DCA_ADV_STATIC_MAX_N' in lib.dca_last_error(None)
