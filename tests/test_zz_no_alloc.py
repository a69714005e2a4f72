"""b200ctc_forced_align_static: the allocation-free variant of b200ctc_forced_align (scratch in a
file-scope `static __device__` array, include/b200ctc.h).  Collected LAST on purpose (file name):
a failure here cannot hide another test.

Checked through the C ABI against the existing workspace entry point, integer-identical (the
tolerance of the other forced-alignment tests), on two calls in a row with DIFFERENT shapes - the
second, smaller problem runs over the first one's stale scratch contents."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_BYTES = 4 << 20  # B200CTC_ALIGN_STATIC_MAX_BYTES


def _case(seed, T, V, L, dev):
    g = torch.Generator().manual_seed(seed)
    lp = torch.log_softmax(torch.randn(T, V, generator=g) * 2.0, dim=-1).to(dev).contiguous()
    tgt = torch.randint(1, V, (L,), generator=g)
    tgt[L // 2] = tgt[L // 2 - 1]  # one repeated label: the skip transition must be refused there
    return lp, tgt.to(dev).contiguous()


def test_static_scratch_matches_workspace_entry_point_twice():
    from espnet1_b200 import _lib
    from espnet1_b200._lib import check, ptr, stream_ptr

    L = _lib.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    st = stream_ptr(dev)
    # every buffer exists BEFORE the calls under test: they must not allocate
    cases = []
    for seed, T, V, tl in ((11, 700, 96, 120), (12, 231, 40, 33)):
        lp, tgt = _case(seed, T, V, tl, dev)
        n = L.b200ctc_align_workspace_bytes(T, tl)
        assert T * (2 * tl + 1) * 4 <= MAX_BYTES
        ws = torch.empty(n, dtype=torch.uint8, device=dev)
        ref = torch.full((T,), -7, dtype=torch.int64, device=dev)
        out = torch.full((T,), -9, dtype=torch.int64, device=dev)
        cases.append((lp, tgt, T, V, tl, ws, n, ref, out))
    for lp, tgt, T, V, tl, ws, n, ref, out in cases:
        check(L.b200ctc_forced_align(ptr(lp), ptr(tgt), T, V, tl, 0, ptr(ref), ptr(ws), n, st), "forced_align")
    torch.cuda.synchronize(dev)
    free0, _ = torch.cuda.mem_get_info(dev)
    stats0 = torch.cuda.memory_stats(dev).get("allocation.all.allocated", 0)
    for lp, tgt, T, V, tl, ws, n, ref, out in cases:  # two calls in a row, no sync in between
        check(L.b200ctc_forced_align_static(ptr(lp), ptr(tgt), T, V, tl, 0, ptr(out), st), "forced_align_static")
    torch.cuda.synchronize(dev)
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, "the allocation-free entry point changed the device's free memory"
    assert torch.cuda.memory_stats(dev).get("allocation.all.allocated", 0) == stats0
    for lp, tgt, T, V, tl, ws, n, ref, out in cases:
        r, o = ref.cpu().numpy(), out.cpu().numpy()
        assert (r >= 0).all()
        np.testing.assert_array_equal(o, r)
        # every frame emits the blank or one of the target's labels.  (Collapsing the path need
        # not give the target back on unpeaked random inputs: the reference lets state 0 take its
        # predecessor from index -1 = the LAST state, espnet/nets/pytorch_backend/ctc.py:163-226,
        # and the kernel reproduces that wrap - the best path here goes round twice.)
        assert np.isin(o, np.concatenate([[0], tgt.cpu().numpy()])).all()
    # and once more in the opposite order (small problem first, then the large one)
    for lp, tgt, T, V, tl, ws, n, ref, out in reversed(cases):
        out.fill_(-9)
        check(L.b200ctc_forced_align_static(ptr(lp), ptr(tgt), T, V, tl, 0, ptr(out), st), "forced_align_static")
    torch.cuda.synchronize(dev)
    for lp, tgt, T, V, tl, ws, n, ref, out in cases:
        np.testing.assert_array_equal(out.cpu().numpy(), ref.cpu().numpy())


def test_static_scratch_rejects_sizes_beyond_the_maximum():
    from espnet1_b200 import _lib
    from espnet1_b200._lib import ptr, stream_ptr

    L = _lib.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    T, V, tl = 4000, 8, 200  # 4000 * 401 * 4 = 6.4 MB > 4 MiB
    assert T * (2 * tl + 1) * 4 > MAX_BYTES
    lp = torch.zeros(T, V, device=dev)
    tgt = torch.ones(tl, dtype=torch.int64, device=dev)
    out = torch.full((T,), -9, dtype=torch.int64, device=dev)
    rc = L.b200ctc_forced_align_static(ptr(lp), ptr(tgt), T, V, tl, 0, ptr(out), stream_ptr(dev))
    assert rc == -3  # B200CTC_ERR_WORKSPACE
    msg = L.b200ctc_last_error().decode()
    assert "B200CTC_ALIGN_STATIC_MAX_BYTES" in msg and "b200ctc_forced_align" in msg
    torch.cuda.synchronize(dev)
    assert int((out != -9).sum()) == 0  # nothing was launched
    rc = L.b200ctc_forced_align_static(None, ptr(tgt), 8, V, 2, 0, ptr(out), stream_ptr(dev))
    assert rc == -1 and b"null pointer" in L.b200ctc_last_error()
    assert isinstance(ctypes.c_int(rc).value, int)
