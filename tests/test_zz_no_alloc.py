"""dcd_seq_encode_noalloc (include/docodo_b200.h): the allocation-free variant of the list encoder, checked through the C-ABI
against the existing entry point dcd_seq_encode on two calls in a row (the scratch array keeps the previous call's contents).
Integer work: the streams must be identical ushort for ushort.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import docodo_b200 as dcd
from docodo_b200.capi import DocodoError, check, lib

pytestmark = pytest.mark.gpu

MAX_VALUES = 1 << 24   # DCD_NOALLOC_MAX_VALUES


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def encode_existing(ix, values):
    out = np.zeros(max(5 * len(values), 1), dtype=np.uint16)
    n = C.c_uint64()
    check(ix._ctx, lib().dcd_seq_encode(ix._ctx, _ptr(values), len(values), _ptr(out), len(out), C.byref(n)))
    return out[:n.value].copy()


def encode_noalloc(ix, values, cap=None):
    import torch
    dev = torch.device("cuda:0")
    d_vals = torch.from_numpy(values.view(np.int64)).to(dev)
    cap = max(5 * len(values), 1) if cap is None else cap
    d_out = torch.zeros(max(cap, 1), dtype=torch.int16, device=dev)
    d_n = torch.full((2,), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    check(ix._ctx, lib().dcd_seq_encode_noalloc(ix._ctx, C.c_void_p(d_vals.data_ptr()), len(values), C.c_void_p(d_out.data_ptr()), cap,
                                                C.c_void_p(d_n.data_ptr())))
    torch.cuda.synchronize()   # the call only enqueues on the context's stream
    n, flag = (int(x) for x in d_n.cpu().numpy())
    return d_out.cpu().numpy().view(np.uint16)[:n].copy(), flag


def ascending(rng, n, max_gap):
    gaps = rng.integers(0, max_gap, size=n, dtype=np.uint64)
    gaps[rng.integers(0, n, size=min(n, 100))] += np.uint64(1 << 31)   # some three-ushort differences (the list stays below 2^40)
    return np.cumsum(gaps, dtype=np.uint64)


def test_noalloc_encode_matches_existing_entry_point_twice():
    rng = np.random.default_rng(20261105)
    with dcd.Index() as ix:
        # two calls in a row with different lengths: the second one finds the first one's status words and ticket in the scratch
        for n, gap in ((300_001, 70_000), (41_000, 300), (2048, 5), (1, 9)):
            v = ascending(rng, n, gap)
            assert int(v[-1]) < (1 << 40)
            want = encode_existing(ix, v)
            got, flag = encode_noalloc(ix, v)
            assert flag == 0
            assert got.size == want.size
            assert np.array_equal(got, want)
        # the same input again: identical stream
        again, flag = encode_noalloc(ix, v)
        assert flag == 0 and np.array_equal(again, want)


def test_noalloc_encode_limits_and_flags():
    rng = np.random.default_rng(7)
    with dcd.Index() as ix:
        v = ascending(rng, 5000, 1000)
        want = encode_existing(ix, v)
        # output too small: flagged, nothing written beyond the capacity
        got, flag = encode_noalloc(ix, v, cap=want.size - 1)
        assert flag == 1
        # not ascending: flagged
        w = v.copy()
        w[100], w[101] = w[101], w[100]
        if w[100] != w[101]:
            _, flag = encode_noalloc(ix, w)
            assert flag == 2
        # empty list
        got, flag = encode_noalloc(ix, np.zeros(0, dtype=np.uint64))
        assert got.size == 0 and flag == 0
        # a size beyond the fixed scratch is refused with a clear error and nothing is enqueued
        import torch
        d_n = torch.zeros(2, dtype=torch.int64, device="cuda:0")
        st = lib().dcd_seq_encode_noalloc(ix._ctx, C.c_void_p(d_n.data_ptr()), MAX_VALUES + 1, C.c_void_p(d_n.data_ptr()), 1,
                                          C.c_void_p(d_n.data_ptr()))
        assert st == -4   # DCD_E_CAPACITY
        with pytest.raises(DocodoError, match="DCD_NOALLOC_MAX_VALUES"):
            check(ix._ctx, st)
        # and the entry point still works afterwards
        got, flag = encode_noalloc(ix, v)
        assert flag == 0 and np.array_equal(got, want)


def test_noalloc_encode_largest_size():
    """n == DCD_NOALLOC_MAX_VALUES uses every status word of the scratch array; checked by decoding (running sum of the
    differences) instead of against the oracle: the stream must reproduce the values."""
    rng = np.random.default_rng(11)
    with dcd.Index() as ix:
        v = np.cumsum(rng.integers(1, 30000, size=MAX_VALUES, dtype=np.uint64), dtype=np.uint64)
        got, flag = encode_noalloc(ix, v)
        assert flag == 0
        g = got.astype(np.uint64)
        ends = (g & np.uint64(0x8000)) == 0
        assert int(ends.sum()) == MAX_VALUES
        # every difference here is below 2^15 (one ushort) -> the stream is the differences themselves
        assert got.size == MAX_VALUES
        assert np.array_equal(np.cumsum(g, dtype=np.uint64), v)
