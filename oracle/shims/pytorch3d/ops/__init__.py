import torch
from oracle import tp_ops as _t


def sample_farthest_points(points, lengths=None, K=50, random_start_point=False):
    """API of pytorch3d.ops.sample_farthest_points; the random start is drawn exactly as
    upstream does: one torch.randint(high=len, size=(1,)).item() per cloud, in order."""
    N, P, _ = points.shape
    if lengths is None:
        lengths = torch.full((N,), P, dtype=torch.int64)
    if isinstance(K, int):
        K = torch.full((N,), K, dtype=torch.int64)
    K = torch.as_tensor(K).to(torch.int64)
    starts = []
    for n in range(N):
        starts.append(int(torch.randint(high=int(lengths[n]), size=(1,)).item()) if random_start_point else 0)
    idx = _t.sample_farthest_points(points, lengths, K, starts)
    safe = idx.clamp(min=0)
    pts = torch.gather(points, 1, safe[:, :, None].expand(-1, -1, points.shape[2]))
    return pts, idx


def ball_query(p1, p2, lengths1=None, lengths2=None, K=500, radius=0.2, return_nn=True):
    idx = _t.ball_query(p1, p2, lengths1, lengths2, K, radius)
    return None, idx, None


def knn_points(p1, p2, lengths1=None, lengths2=None, norm=2, K=1, version=-1, return_nn=False, return_sorted=True):
    d, idx = _t.knn_points(p1, p2, lengths1, lengths2, K)
    return d, idx, None
