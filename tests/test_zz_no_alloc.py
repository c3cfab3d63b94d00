"""ccn_fps_static: the farthest-point sampler with its scratch memory in a file-scope device array of the library (no
cudaMalloc anywhere on the call path: callers under CUDA graph capture, serving loops, harnesses that forbid allocation).
It has no counterpart in the reference; it must return exactly what ccn_fps returns (index work: bit-exact), on two calls in a
row that SHARE the array - the second call on a smaller, different input, so stale contents of the first would show.
Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_POINTS = 1 << 18            # CCN_FPS_STATIC_MAX_POINTS (include/ccn_b200.h)
CCN_ESCRATCH_LIMIT = -6


def _clouds(lengths, seed, kmax=None):
    gen = torch.Generator().manual_seed(seed)
    n = sum(lengths)
    pos = torch.rand(n, 3, generator=gen)
    # duplicated points: the lowest-index tie rule decides
    if n > 40:
        pos[17] = pos[3]
        pos[n - 1] = pos[n - 9]
    sptr = torch.zeros(len(lengths) + 1, dtype=torch.int32)
    sptr[1:] = torch.cumsum(torch.tensor(lengths), 0)
    ksel = torch.tensor([min(max(1, (l * 3 + 9) // 10), kmax or l) for l in lengths], dtype=torch.int32)
    start = torch.tensor([int(torch.randint(0, l, (1,), generator=gen)) for l in lengths], dtype=torch.int32)
    return pos.cuda(), sptr.cuda(), ksel.cuda(), start.cuda(), int(ksel.sum()), max(lengths)


def _both(lengths, seed, kmax=None):
    from curvecloudnet_b200.native import call
    pos, sptr, ksel, start, total, max_len = _clouds(lengths, seed, kmax)
    n, B = pos.shape[0], len(lengths)
    # the existing entry point: scratch supplied by the caller
    closest = torch.empty(n, dtype=torch.float32, device="cuda")
    sel = torch.empty(n, dtype=torch.int32, device="cuda")
    want = torch.full((total,), -7, dtype=torch.int64, device="cuda")
    want_total = torch.zeros(1, dtype=torch.int32, device="cuda")
    call("ccn_fps", pos, sptr, B, n, ksel, start, max_len, closest, sel, want, want_total)
    # the no-allocation entry point: inputs and outputs only
    got = torch.full((total,), -7, dtype=torch.int64, device="cuda")
    got_total = torch.zeros(1, dtype=torch.int32, device="cuda")
    call("ccn_fps_static", pos, sptr, B, n, ksel, start, max_len, got, got_total)
    torch.cuda.synchronize()
    return got, got_total, want, want_total, total


def test_fps_static_equals_fps_on_two_calls_in_a_row():
    # call 1: a 20 k-point cloud (the global-memory `closest` variant, > 8192 points) next to small ones;
    # call 2: fewer, shorter clouds - every scratch entry it touches holds call 1's values unless the entry point clears them
    for lengths, seed in (([20000, 700, 3000, 1], 1), ([900, 2500], 2), ([5000], 3)):
        got, got_total, want, want_total, total = _both(lengths, seed)
        assert int(want_total) == total and int(got_total) == total
        assert torch.equal(got, want), "ccn_fps_static picked other points than ccn_fps (%s)" % (lengths,)
        assert bool((got[1:] > got[:-1]).all()), "selected indices must be sorted ascending, without repeats"


def test_fps_static_rejects_sizes_beyond_its_scratch():
    from curvecloudnet_b200.native import CcnError, call
    n = MAX_POINTS + 1
    pos = torch.zeros(n, 3, device="cuda")
    sptr = torch.tensor([0, n], dtype=torch.int32, device="cuda")
    ksel = torch.tensor([4], dtype=torch.int32, device="cuda")
    start = torch.tensor([0], dtype=torch.int32, device="cuda")
    out = torch.full((4,), -7, dtype=torch.int64, device="cuda")
    tot = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    with pytest.raises(CcnError) as ei:
        call("ccn_fps_static", pos, sptr, 1, n, ksel, start, n, out, tot)
    assert ei.value.code == CCN_ESCRATCH_LIMIT
    torch.cuda.synchronize()
    assert int(tot) == -7 and bool((out == -7).all()), "a rejected call must not launch anything"
    # the limit itself is accepted
    got, got_total, want, want_total, total = _both([MAX_POINTS - 5000, 5000], 4, kmax=200)
    assert torch.equal(got, want) and int(got_total) == total
