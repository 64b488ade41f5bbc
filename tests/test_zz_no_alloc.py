"""pbhc_compact_ids_static - the allocation-free variant of pbhc_compact_ids (scratch in a file-scope device array of the
library; no counterpart in the reference) - against pbhc_compact_ids through
the C ABI, on two calls in a row that share the array.  Integer output: exact equality.  This file is collected last on
purpose."""
import ctypes as C

import pytest
import torch

from pbhc_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _both(L, flags, stream):
    n = flags.numel()
    out = []
    for static in (False, True):
        ids = torch.full((n,), -7, device=DEV, dtype=torch.int32)
        count = torch.full((1,), -1, device=DEV, dtype=torch.int32)
        if static:
            rc = L.pbhc_compact_ids_static(_lib.ptr(flags), n, _lib.ptr(ids), _lib.ptr(count), stream)
        else:
            scratch = torch.zeros(max(int(L.pbhc_compact_scratch_ints(n)), 1), device=DEV, dtype=torch.int32)
            rc = L.pbhc_compact_ids(_lib.ptr(flags), n, _lib.ptr(ids), _lib.ptr(count), _lib.ptr(scratch), stream)
        _lib.check(rc)
        out.append((ids, count))
    return out


def test_static_scratch_variant_matches_caller_scratch_variant_twice_in_a_row():
    L = _lib.lib()
    stream = _lib.stream_ptr(torch.device(DEV))
    gen = torch.Generator().manual_seed(11)
    # two calls in a row on the shared array: a large dense mask first, then a smaller sparse one whose block counts must
    # not see what the first call left behind; then the first again
    big = ((torch.rand(300007, generator=gen) < 0.6).to(torch.int64) * 3).to(DEV)
    small = (torch.rand(65536, generator=gen) < 0.02).to(torch.int64).to(DEV)
    for flags in (big, small, big, small[:37].contiguous()):
        (ids_a, cnt_a), (ids_b, cnt_b) = _both(L, flags, stream)
        torch.cuda.synchronize()
        want = flags.nonzero().flatten().to(torch.int32)
        assert int(cnt_b.item()) == int(cnt_a.item()) == want.numel()
        assert torch.equal(ids_b, ids_a)  # every entry, the untouched tail included
        assert torch.equal(ids_b[:want.numel()], want)


def test_static_scratch_variant_rejects_sizes_beyond_its_array():
    L = _lib.lib()
    flags = torch.zeros(16, device=DEV, dtype=torch.int64)
    ids = torch.zeros(16, device=DEV, dtype=torch.int32)
    count = torch.zeros(1, device=DEV, dtype=torch.int32)
    too_many = _lib.COMPACT_STATIC_MAX_N + 1
    assert L.pbhc_compact_ids_static(_lib.ptr(flags), too_many, _lib.ptr(ids), _lib.ptr(count), None) == -3  # PBHC_E_RANGE, before any launch
    assert b"PBHC_COMPACT_STATIC_MAX_N" in L.pbhc_last_error()
    assert L.pbhc_compact_ids_static(None, 16, _lib.ptr(ids), _lib.ptr(count), None) == -1
    assert L.pbhc_compact_ids_static(_lib.ptr(flags), 16, _lib.ptr(ids), _lib.ptr(count), None) == 0
    torch.cuda.synchronize()
    assert int(count.item()) == 0
