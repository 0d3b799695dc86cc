"""rt_sort_pairs_noalloc: the allocation-free variant of the one-sweep radix sort (include/rt_b200.h).

It has no counterpart in the reference (added for callers that cannot
call cudaMalloc).  Its scratch is a fixed file-scope device array shared by every call, so the checks are: the
output equals the existing allocating entry point's (rt_debug_sort_pairs) bit for bit - integer work, no
tolerance - on two calls in a row with different sizes and contents (the second call meets the first call's
leftovers in the scratch), sizes beyond RT_SORT_NOALLOC_MAX_PAIRS are refused with a clear error, and the call
itself allocates no device memory.  Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def reference_sort(rt, keys, vals):
    """the existing entry point: host pointers, allocates its own device buffers"""
    ko, vo = np.empty_like(keys), np.empty_like(vals)
    rc = rt.rt_lib().rt_debug_sort_pairs(0, keys.size, keys.ctypes.data, vals.ctypes.data, ko.ctypes.data, vo.ctypes.data)
    assert rc == 0, rt.rt_lib().rt_last_error().decode()
    return ko, vo


def noalloc_sort(rt, torch, keys, vals):
    """the new entry point: the caller's device buffers, sorted in place on the caller's stream"""
    dk = torch.from_numpy(keys.view(np.int64)).to("cuda:0")
    dv = torch.from_numpy(vals.view(np.int32)).to("cuda:0")
    torch.cuda.synchronize()
    before = torch.cuda.mem_get_info(0)[0]
    rc = rt.rt_lib().rt_sort_pairs_noalloc(0, keys.size, C.c_void_p(dk.data_ptr()), C.c_void_p(dv.data_ptr()),
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, rt.rt_lib().rt_last_error().decode()
    torch.cuda.synchronize()
    after = torch.cuda.mem_get_info(0)[0]
    return dk.cpu().numpy().view(np.uint64), dv.cpu().numpy().view(np.uint32), before - after


def test_noalloc_sort_equals_the_allocating_sort_on_two_calls_in_a_row(gpu_rt):
    import torch
    rt = gpu_rt
    rng = np.random.default_rng(20251105)
    cases = []
    # call 1: the largest size the fixed scratch holds, full 64-bit keys
    n1 = rt.SORT_NOALLOC_MAX_PAIRS
    cases.append((rng.integers(0, 2 ** 64, n1, dtype=np.uint64), np.arange(n1, dtype=np.uint32)))
    # call 2, straight after: a smaller, ragged size (not a multiple of the 4096-key tile), many duplicate keys so
    # that stability shows, values that are not a permutation
    n2 = 300_001
    cases.append((rng.integers(0, 1000, n2, dtype=np.uint64) << np.uint64(37), rng.integers(0, 2 ** 32, n2, dtype=np.uint32)))
    # call 3: one pair, then call 4: the full size again with other contents
    cases.append((np.array([42], np.uint64), np.array([7], np.uint32)))
    cases.append((rng.integers(0, 2 ** 63, n1, dtype=np.uint64), rng.integers(0, 2 ** 32, n1, dtype=np.uint32)))
    first = True
    for keys, vals in cases:
        ko, vo, grew = noalloc_sort(rt, torch, keys, vals)
        rk, rv = reference_sort(rt, keys, vals)
        assert np.array_equal(ko, rk) and np.array_equal(vo, rv), keys.size
        order = np.argsort(keys, kind="stable")      # and both equal a stable host sort
        assert np.array_equal(ko, keys[order]) and np.array_equal(vo, vals[order])
        if not first:   # (the very first CUDA call of a process may load the module lazily)
            assert grew <= 0, f"device memory in use grew by {grew} bytes during rt_sort_pairs_noalloc"
        first = False


def test_noalloc_sort_refuses_sizes_beyond_its_scratch(gpu_rt):
    import torch
    rt = gpu_rt
    n = rt.SORT_NOALLOC_MAX_PAIRS + 1
    dk = torch.zeros(n, dtype=torch.int64, device="cuda:0")
    dv = torch.arange(n, dtype=torch.int32, device="cuda:0")
    rc = rt.rt_lib().rt_sort_pairs_noalloc(0, n, C.c_void_p(dk.data_ptr()), C.c_void_p(dv.data_ptr()), None)
    msg = rt.rt_lib().rt_last_error().decode()
    assert rc == rt.ERR_UNSUPPORTED and "RT_SORT_NOALLOC_MAX_PAIRS" in msg and str(rt.SORT_NOALLOC_MAX_PAIRS) in msg, (rc, msg)
    torch.cuda.synchronize()
    assert torch.equal(dv, torch.arange(n, dtype=torch.int32, device="cuda:0"))      # nothing was touched
    # n = 0 is a no-op, null pointers are an argument error
    assert rt.rt_lib().rt_sort_pairs_noalloc(0, 0, None, None, None) == 0
    assert rt.rt_lib().rt_sort_pairs_noalloc(0, 5, None, None, None) == rt.ERR_INVALID_ARGUMENT
