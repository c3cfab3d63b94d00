"""Host logic of autograd.py (no GPU): activation planes / column sums a producer attaches to a tensor are only trusted while the
tensor still holds the values they were computed from (an in-place accumulation into the same TensorImpl - a second
gradient contribution - must invalidate them)."""
import torch


class _FakePlanes:
    def __init__(self, shape, n_planes):
        self.shape, self.n_planes = tuple(shape), n_planes


def test_attached_planes_are_dropped_after_an_in_place_write():
    from curvecloudnet_b200 import autograd as A
    t = torch.zeros(5, 16)
    t._ccn_planes = _FakePlanes(t.shape, 2)
    t._ccn_colsum = torch.ones(16)
    assert A._planes_of(t, 2) is None and A._colsum_of(t) is None        # never stamped: not trusted
    A._tag(t)
    assert A._planes_of(t, 2) is t._ccn_planes and A._colsum_of(t) is t._ccn_colsum
    assert A._planes_of(t, 1) is None                                      # another precision mode
    t.add_(1.0)                                                            # what the engine's InputBuffer does on a second contribution
    assert A._planes_of(t, 2) is None and A._colsum_of(t) is None
