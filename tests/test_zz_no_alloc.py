"""pdgs_sort_pairs_static: the pair sort for hosts that may not allocate device memory.

Who needs it: programs in which this library is a guest and the device memory is not theirs to ask for -- an
engine or serving process whose own memory manager claimed the whole budget at start-up, a plugin that runs inside
another program's frame loop (a cudaMalloc there synchronises the whole device and stalls every other stream), a
deployment with a fixed memory map.  Such a caller still owns its input and output buffers; what it cannot supply is
a fresh scratch buffer per call.  The entry point therefore keeps the sort's working memory in a fixed array inside
libpdgs.so (about 2.7 MiB per GPU, good for up to PDGS_SORT_STATIC_MAX_PAIRS = 2^20 pairs).

What is checked here, through the C ABI only:
  * the output equals pdgs_sort_pairs' output bit for bit (integer work: no tolerance), on several calls in a
    row that share the array -- a large sort followed by smaller ones, so anything a call failed to clear would
    show up in the next one;
  * sizes beyond the limit are refused with an error that names the limit, before any GPU work.

This file is named so that pytest collects it last: with `-x` a failure here must not hide the other tests.
"""
import ctypes as C

import pytest
import torch

from pdgs_b200 import _lib as L

MAX_PAIRS = 1 << 20  # PDGS_SORT_STATIC_MAX_PAIRS in include/pdgs.h


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _pairs(n, end_bit, seed, dev):
    gen = torch.Generator().manual_seed(seed)
    hi = 2 ** min(end_bit, 62)
    keys = torch.randint(0, hi, (n,), generator=gen, dtype=torch.int64)
    if n > 16:  # runs of equal keys: the sort has to be stable
        keys[n // 3: n // 3 + n // 8] = keys[n // 3]
    vals = torch.randperm(n, generator=gen).to(torch.int32)
    return keys.to(dev), vals.to(dev)


def _sort_with_caller_scratch(lib, keys, vals, end_bit, stream):
    """The existing entry point, pdgs_sort_pairs, with a scratch buffer the caller allocates."""
    n = keys.numel()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    nb = lib.pdgs_sort_scratch_bytes(n)
    scratch = torch.empty(nb, dtype=torch.uint8, device=keys.device)
    L.check(lib.pdgs_sort_pairs(_ptr(keys), _ptr(vals), _ptr(ko), _ptr(vo), n, end_bit, _ptr(scratch), nb, stream),
            "pdgs_sort_pairs")
    return ko, vo


def _sort_with_builtin_scratch(lib, keys, vals, end_bit, stream):
    """The entry point under test: inputs and outputs only, no scratch argument."""
    n = keys.numel()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    L.check(lib.pdgs_sort_pairs_static(_ptr(keys), _ptr(vals), _ptr(ko), _ptr(vo), n, end_bit, stream),
            "pdgs_sort_pairs_static")
    return ko, vo


@pytest.mark.gpu
def test_builtin_scratch_sort_equals_caller_scratch_sort_on_consecutive_calls():
    lib = L.lib()
    dev = torch.device("cuda:0")
    with torch.cuda.device(dev):
        stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        # (n, end_bit): the largest accepted size first, then smaller ones that reuse the array it left behind;
        # odd and even numbers of 8-bit passes (the two ping-pong arrangements), a ragged tail, a single pair
        calls = [(MAX_PAIRS, 45), (70001, 40), (70001, 40), (3072, 64), (4097, 17), (1, 8), (MAX_PAIRS, 33)]
        for i, (n, end_bit) in enumerate(calls):
            keys, vals = _pairs(n, end_bit, seed=100 + i, dev=dev)
            want_k, want_v = _sort_with_caller_scratch(lib, keys.clone(), vals.clone(), end_bit, stream)
            got_k, got_v = _sort_with_builtin_scratch(lib, keys.clone(), vals.clone(), end_bit, stream)
            torch.cuda.synchronize(dev)
            assert torch.equal(got_k, want_k), f"keys differ on call {i} (n={n}, end_bit={end_bit})"
            assert torch.equal(got_v, want_v), f"values differ on call {i} (n={n}, end_bit={end_bit})"
            # and both are what a stable sort over the low end_bit bits gives
            mask = (1 << end_bit) - 1 if end_bit < 63 else -1
            ref_k, idx = torch.sort((keys & mask).cpu(), stable=True)
            assert torch.equal((got_k & mask).cpu(), ref_k)
            assert torch.equal(got_v.cpu(), vals.cpu()[idx])
        # n == 0 is accepted and does nothing
        e = torch.empty(0, dtype=torch.int64, device=dev)
        ev = torch.empty(0, dtype=torch.int32, device=dev)
        assert lib.pdgs_sort_pairs_static(_ptr(e), _ptr(ev), _ptr(e), _ptr(ev), 0, 32, stream) == L.PDGS_OK


def test_builtin_scratch_sort_refuses_sizes_beyond_its_limit():
    """Refused on the host, before any CUDA call (so this runs without a GPU too): dummy non-NULL pointers."""
    lib = L.lib()
    dummy = C.c_void_p(256)
    rc = lib.pdgs_sort_pairs_static(dummy, dummy, dummy, dummy, MAX_PAIRS + 1, 32, None)
    assert rc == L.PDGS_E_INVALID
    msg = L.last_error()
    assert str(MAX_PAIRS) in msg and "pdgs_sort_pairs" in msg
    assert lib.pdgs_sort_pairs_static(dummy, dummy, dummy, dummy, -1, 32, None) == L.PDGS_E_INVALID
    assert lib.pdgs_sort_pairs_static(dummy, dummy, dummy, dummy, 10, 65, None) == L.PDGS_E_INVALID
    assert lib.pdgs_sort_pairs_static(None, dummy, dummy, dummy, 10, 32, None) == L.PDGS_E_INVALID
