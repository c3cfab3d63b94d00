"""Loader for the committed fixtures made by oracle/make_golden.py (int32 -> int64 widening)."""
import os

import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _widen(obj):
    if torch.is_tensor(obj):
        return obj.to(torch.int64) if obj.dtype == torch.int32 else obj
    if isinstance(obj, dict):
        return {k: _widen(v) for k, v in obj.items()}
    return obj


def load(name):
    return _widen(torch.load(os.path.join(GOLDEN, name), weights_only=False))


class FixedDraws:
    """Draws object that replays recorded values (same interface as oracle.ccn_point_ops.Draws)."""

    def __init__(self, u=None, noise=None, starts=None):
        self.u, self.noise, self.starts = u, noise, list(starts) if starts is not None else None

    def curve_phase(self):
        return self.u

    def voxel_noise(self, n):
        assert self.noise.shape[0] == n
        return self.noise

    def fps_start(self, length):
        return int(self.starts.pop(0))


def rel_err(a, b):
    """max |a-b| / max |b|  (the 'relative' of BASELINE.json's 1e-4 / 2e-2 tolerances)."""
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp(min=1e-30))
