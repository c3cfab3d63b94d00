from oracle.tp_nn import MLP  # noqa
from . import conv, glob, inits  # noqa
def fps(*a, **k):
    raise NotImplementedError("torch_geometric.nn.fps: only used by the dead SAModuleSlow (pointnet2.py:21)")
def radius(*a, **k):
    raise NotImplementedError("torch_geometric.nn.radius: only used by the dead SAModuleSlow (pointnet2.py:22)")
def knn(x, y, k, batch_x=None, batch_y=None, cosine=False, num_workers=1):
    """torch_geometric.nn.knn -> torch_cluster.knn (1.6): for every row of y the k nearest rows of x of the same batch id,
    by squared Euclidean distance in the tensors' dtype; -> LongTensor (2, E): [y index, x index], grouped by y, nearest first,
    fewer than k pairs where the batch has fewer rows.  Published semantics restated by brute force (the package is neither
    installed nor vendored); only the scanning simulator's curvature estimate calls it (curvature_approx.py:29)."""
    import torch
    assert not cosine
    if batch_x is None:
        batch_x = torch.zeros(x.shape[0], dtype=torch.long)
    if batch_y is None:
        batch_y = torch.zeros(y.shape[0], dtype=torch.long)
    rows, cols = [], []
    for b in torch.unique(batch_y).tolist():
        iy = torch.nonzero(batch_y == b).flatten()
        ix = torch.nonzero(batch_x == b).flatten()
        if ix.numel() == 0:
            continue
        d = ((y[iy].unsqueeze(1) - x[ix].unsqueeze(0)) ** 2).sum(-1)
        kk = min(k, ix.numel())
        nn = torch.topk(d, kk, dim=1, largest=False, sorted=True).indices
        rows.append(iy.unsqueeze(1).expand(-1, kk).reshape(-1))
        cols.append(ix[nn].reshape(-1))
    return torch.stack([torch.cat(rows), torch.cat(cols)], dim=0)
