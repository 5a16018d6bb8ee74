"""-m gpu, collected last on purpose: mg_invalid_readback_noalloc (include/mcaat_gpu.h; no counterpart in the
reference) against mg_invalid_readback, through the C ABI.

The no-allocation entry stages the validity words in a static device array of the library instead of a pool
allocation, so (a) its words must equal the allocating entry's bit for bit — the tolerance of every byte/bit
parity test here is zero — on consecutive calls, including after the bitmap changed in between (the array keeps
the last call's contents), and (b) graphs beyond MG_NOALLOC_MAX_ITEMS are refused with MG_ELIMIT.
"""
import ctypes as C

import numpy as np
import pytest

import golden_util as G
from mcaat_b200 import DeviceGraph, gpu

pytestmark = pytest.mark.gpu

MG_NOALLOC_MAX_ITEMS = 1 << 25
MG_ELIMIT = -4


def _noalloc_words(g):
    # caller-owned output, pre-filled with a pattern so that words the call does not write are seen
    out = np.full((g.n + 63) // 64, 0xA5A5A5A5A5A5A5A5, np.uint64)
    rc = gpu.lib().mg_invalid_readback_noalloc(g.h, out.ctypes.data)
    assert rc == 0, gpu.lib().mg_last_error()
    return out


def test_no_alloc_readback_equals_the_allocating_entry_on_consecutive_calls():
    big = DeviceGraph(G.pack_raw(G.load_graph_npz("strain")))
    small = DeviceGraph(G.pack_raw(G.load_graph_npz("tiny")))
    assert small.n < big.n <= MG_NOALLOC_MAX_ITEMS
    launches = gpu.launch_count()
    # call 1: the pristine bitmap
    want = big.invalid_words()
    got = _noalloc_words(big)
    assert got.dtype == want.dtype and np.array_equal(got, want)
    assert gpu.launch_count() >= launches + 2  # both entries launched their gather kernel
    # call 2, directly after: a smaller graph — the array still holds the larger graph's words beyond its end
    assert np.array_equal(_noalloc_words(small), small.invalid_words())
    # call 3: the first graph again after pruning changed its bitmap (stale contents must not leak through)
    big.prune()
    after = big.invalid_words()
    assert not np.array_equal(after, want)
    assert np.array_equal(_noalloc_words(big), after)
    assert np.array_equal(_noalloc_words(big), after)  # and once more, unchanged
    # the Python mirror of the entry
    assert np.array_equal(big.invalid_words_noalloc(), after)
    big.reset()
    assert np.array_equal(_noalloc_words(big), want)


def test_no_alloc_readback_rejects_bad_arguments_and_sizes_beyond_the_static_scratch():
    L = gpu.lib()
    g = DeviceGraph(G.pack_raw(G.load_graph_npz("tiny")))
    assert L.mg_invalid_readback_noalloc(None, None) == -1
    assert L.mg_invalid_readback_noalloc(g.h, None) == -1
    # a graph one validity block past the limit: only its size matters (it is built and never searched)
    n = MG_NOALLOC_MAX_ITEMS + 256
    nib = np.tile(np.array([1, 2, 3, 4], np.uint8), n // 4)
    w = (nib[0::2] | (nib[1::2] << 4)).astype(np.uint8).view("<u8").copy()
    ones = np.full((n + 63) // 64, np.uint64(0xFFFFFFFFFFFFFFFF), np.uint64)
    packed = dict(n=n, k=21, w=w, last=ones, tip=np.zeros_like(ones), mult=np.ones(n, np.uint16),
                  f=np.array([-1, 0, n // 4, n // 2, 3 * n // 4, n], np.int64))
    huge = DeviceGraph(packed)
    out = np.zeros((n + 63) // 64, np.uint64)
    launches = gpu.launch_count()
    assert L.mg_invalid_readback_noalloc(huge.h, out.ctypes.data) == MG_ELIMIT
    msg = L.mg_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "MG_NOALLOC_MAX_ITEMS" in msg
    assert gpu.launch_count() == launches and not out.any()  # refused before any launch
    assert L.mg_invalid_readback(huge.h, out.ctypes.data) == 0  # the allocating entry has no such limit
    huge.close()
