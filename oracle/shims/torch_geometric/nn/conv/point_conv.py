from oracle.tp_nn import MessagePassing
from ..inits import reset


class PointNetConv(MessagePassing):
    """Constructor/reset contract of torch_geometric.nn.conv.PointNetConv 2.3.0; the
    reference subclass (point_conv.py:12-93) overrides forward/message/aggregate."""

    def __init__(self, local_nn=None, global_nn=None, add_self_loops=True, **kwargs):
        kwargs.setdefault("aggr", "max")
        super().__init__(**kwargs)
        self.local_nn = local_nn
        self.global_nn = global_nn
        self.add_self_loops = add_self_loops
        self.reset_parameters()

    def reset_parameters(self):
        reset(self.local_nn)
        reset(self.global_nn)
