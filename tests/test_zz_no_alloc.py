"""rt_sort_pairs_static: the allocation-free variant of rt_sort_pairs (scratch in a file-scope device array instead of
three cudaMalloc'ed buffers per call; include/rtb_api.h).  Named test_zz_* so that pytest collects it last:
a failure here must not hide the tests before it.

Called through the C-ABI, compared with the existing entry point bit for bit (integer keys and indices: the tolerance of
every other sort test here is exact equality), on calls in a row that share the scratch array -- the second smaller than
the first, so that the first one's leftovers lie behind it -- and with duplicate-heavy keys, where stability shows."""
import ctypes as C

import pytest

import gpu_util as gu
from raytracing_cuda_b200 import _lib

pytestmark = pytest.mark.gpu


def _pairs(torch, n, bits, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    keys = torch.randint(0, 1 << bits, (n,), generator=g, device="cuda", dtype=torch.int32)
    vals = torch.arange(n, device="cuda", dtype=torch.int32)
    return keys, vals


def test_static_scratch_sort_equals_allocating_sort_on_calls_in_a_row():
    import torch
    lib = _lib.load()
    r = gu.renderer()
    h = r._h
    for call, (n, bits) in enumerate(((1_000_003, 30), (100_003, 6), (4_097, 30), (2, 1), (_lib.RT_SORT_STATIC_MAX, 12))):
        keys, vals = _pairs(torch, n, bits, seed=1000 + call)
        k_ref, v_ref, k_new, v_new = keys.clone(), vals.clone(), keys.clone(), vals.clone()
        assert lib.rt_sort_pairs(h, k_ref.data_ptr(), v_ref.data_ptr(), n) == _lib.RT_OK
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = lib.rt_sort_pairs_static(h, k_new.data_ptr(), v_new.data_ptr(), n)
        assert rc == _lib.RT_OK, lib.rt_last_error(h)
        r.synchronize()                                   # the static variant is asynchronous on the context's stream
        assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the device's free memory"
        assert torch.equal(k_new, k_ref) and torch.equal(v_new, v_ref), (call, n)
        ks, order = torch.sort(keys.long(), stable=True)  # and both are the stable ascending sort
        assert torch.equal(k_new.long(), ks) and torch.equal(v_new.long(), order), (call, n)


def test_static_scratch_sort_rejects_what_does_not_fit():
    import torch
    lib = _lib.load()
    r = gu.renderer()
    h = r._h
    keys, vals = _pairs(torch, 64, 8, seed=5)
    before = (keys.clone(), vals.clone())
    rc = lib.rt_sort_pairs_static(h, keys.data_ptr(), vals.data_ptr(), _lib.RT_SORT_STATIC_MAX + 1)   # rejected before any access
    assert rc == _lib.RT_ERR_ARG
    msg = lib.rt_last_error(h).decode()
    assert "exceeds" in msg and str(_lib.RT_SORT_STATIC_MAX) in msg, msg
    r.synchronize()
    assert torch.equal(keys, before[0]) and torch.equal(vals, before[1])
    assert lib.rt_sort_pairs_static(h, None, None, 5) == _lib.RT_ERR_ARG
    assert lib.rt_sort_pairs_static(h, keys.data_ptr(), vals.data_ptr(), 0) == _lib.RT_OK
    assert lib.rt_sort_pairs_static(h, keys.data_ptr(), vals.data_ptr(), -1) == _lib.RT_ERR_ARG
    # and it still works afterwards
    assert lib.rt_sort_pairs_static(h, keys.data_ptr(), vals.data_ptr(), 64) == _lib.RT_OK
    r.synchronize()
    ks, order = torch.sort(before[0].long(), stable=True)
    assert torch.equal(keys.long(), ks) and torch.equal(vals.long(), order)
