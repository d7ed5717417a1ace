"""gmpm_radix_sort_pairs_static_scratch: the sort for callers that may not allocate device memory.

Some hosts of this library get their GPU memory once, at start-up, and are not allowed to call cudaMalloc or
cudaFree afterwards: a simulator inside a real-time loop (either call can stall every stream of the device), or a
plugin inside an application that owns the allocator.  For them the library carries its sort scratch as a fixed
array inside the library image, and the entry point takes device arrays from the caller.  This test drives that entry
point through the C ABI and holds it to the allocating entry point, gmpm_radix_sort_pairs: the two must agree bit for
bit (integer work: the tolerance is zero), call after call, although every call reuses the same array and finds in it
whatever the call before left there.

The file is named so that pytest collects it last: under `pytest -x` a failure here must not hide other tests.  It
passes in a fresh process (the first call does the per-device set-up) and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

MAX_PAIRS = 1 << 20  # GMPM_RADIX_SORT_STATIC_MAX_PAIRS in include/gmpm_c.h


def _header_limit():
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "include", "gmpm_c.h")).read()
    m = re.search(r"#define GMPM_RADIX_SORT_STATIC_MAX_PAIRS \(1u << (\d+)\)", text)
    return 1 << int(m.group(1))


def _static_sort(keys, ids, num_bit, poison=None):
    """keys, ids: numpy uint32.  Device arrays come from torch (the caller's memory, as the entry point expects);
    the library itself allocates nothing.  Returns numpy copies of the outputs."""
    import torch
    from drake_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda:0")
    n = keys.size
    to_dev = lambda a: torch.from_numpy(a.view(np.int32).copy()).to(dev)
    k_in, i_in = to_dev(keys), to_dev(ids)
    k_out = torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev)
    i_out = torch.full((max(n, 1),), -1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream()
        rc = L.gmpm_radix_sort_pairs_static_scratch(k_in.data_ptr(), i_in.data_ptr(), n, int(num_bit),
                                                    k_out.data_ptr(), i_out.data_ptr(), stream.cuda_stream)
        if rc != 0:
            raise _lib.GmpmError(rc, L.gmpm_last_error().decode())
        stream.synchronize()
    # the inputs are read-only for the library
    assert np.array_equal(k_in.cpu().numpy().view(np.uint32), keys)
    assert np.array_equal(i_in.cpu().numpy().view(np.uint32), ids)
    return k_out.cpu().numpy().view(np.uint32)[:n], i_out.cpu().numpy().view(np.uint32)[:n]


def _pairs(n, key_bits, seed):
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 1 << key_bits, size=n, dtype=np.uint32)
    if n >= 3:
        keys[rng.integers(0, n, size=n // 3)] = keys[0]  # many equal keys: the order of ties is part of the result
    ids = rng.permutation(n).astype(np.uint32)
    return keys, ids


def test_limit_constant_matches_header():
    assert _header_limit() == MAX_PAIRS


@pytest.mark.gpu
def test_static_scratch_sort_equals_allocating_sort_on_calls_in_a_row():
    """Several calls back to back, no clean-up in between.  Sizes and bit counts change from call to call so that
    each one finds stale data of a different shape in the shared array: one, two and three digit passes (the
    output is the last ping-pong buffer, so which half of the array is used alternates), a large input followed by
    small ones, the same input twice, and the largest input the array holds."""
    from drake_b200.gpu_mpm import radix_sort_pairs
    calls = [(200003, 21, 21), (4097, 16, 21), (4097, 16, 21), (1, 16, 16), (33, 8, 21), (MAX_PAIRS, 27, 27),
             (100003, 32, 32), (777, 0, 21), (MAX_PAIRS, 21, 21), (31, 21, 21)]
    for c, (n, num_bit, key_bits) in enumerate(calls):
        keys, ids = _pairs(n, key_bits, seed=c)
        want_k, want_i = radix_sort_pairs(keys, ids, num_bit)   # existing entry point: host arrays, cudaMalloc inside
        got_k, got_i = _static_sort(keys, ids, num_bit)
        assert np.array_equal(got_k, want_k), "keys differ in call %d %r" % (c, (n, num_bit))
        assert np.array_equal(got_i, want_i), "ids differ in call %d %r" % (c, (n, num_bit))
        # and it IS a stable sort on the low num_bit bits (independent of the other entry point)
        order = np.argsort(keys & np.uint32((1 << num_bit) - 1 if num_bit < 32 else 0xFFFFFFFF), kind="stable")
        assert np.array_equal(got_i, ids[order]) and np.array_equal(got_k, keys[order])


@pytest.mark.gpu
def test_static_scratch_sort_rejects_what_it_cannot_hold():
    """Beyond the built-in array the call must refuse, say why, and launch nothing; host arrays and overlapping
    arrays are refused too.  The next valid call still works."""
    import torch
    from drake_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda:0")
    n = MAX_PAIRS + 1
    buf = [torch.zeros(n, dtype=torch.int32, device=dev) for _ in range(4)]
    p = [b.data_ptr() for b in buf]
    assert L.gmpm_radix_sort_pairs_static_scratch(p[0], p[1], n, 21, p[2], p[3], None) == -1
    msg = L.gmpm_last_error().decode()
    assert str(MAX_PAIRS) in msg and str(n) in msg
    torch.cuda.synchronize()
    assert all(int(b.abs().max()) == 0 for b in buf)  # untouched
    # overlapping output and input
    assert L.gmpm_radix_sort_pairs_static_scratch(p[0], p[1], 1000, 21, p[0], p[3], None) == -1
    assert "overlap" in L.gmpm_last_error().decode()
    # host arrays
    h = [np.zeros(16, np.uint32) for _ in range(4)]
    hp = [a.ctypes.data_as(C.c_void_p) for a in h]
    assert L.gmpm_radix_sort_pairs_static_scratch(hp[0], hp[1], 16, 21, hp[2], hp[3], None) == -1
    assert "DEVICE" in L.gmpm_last_error().decode()
    # bad bit count; empty input is a no-op
    assert L.gmpm_radix_sort_pairs_static_scratch(p[0], p[1], 16, 33, p[2], p[3], None) == -1
    assert L.gmpm_radix_sort_pairs_static_scratch(None, None, 0, 21, None, None, None) == 0
    keys, ids = _pairs(5000, 21, seed=99)
    from drake_b200.gpu_mpm import radix_sort_pairs
    want = radix_sort_pairs(keys, ids, 21)
    got = _static_sort(keys, ids, 21)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
