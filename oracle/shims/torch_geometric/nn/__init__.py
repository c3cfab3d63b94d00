from oracle.tp_nn import MLP  # noqa
from . import conv, glob, inits  # noqa
def fps(*a, **k):
    raise NotImplementedError("torch_geometric.nn.fps: only used by the dead SAModuleSlow (pointnet2.py:21)")
def radius(*a, **k):
    raise NotImplementedError("torch_geometric.nn.radius: only used by the dead SAModuleSlow (pointnet2.py:22)")
def knn(*a, **k):
    raise NotImplementedError("torch_geometric.nn.knn: imported at point_ops.py:5, never called")
