from oracle.tp_nn import MessagePassing  # noqa
from .point_conv import PointNetConv  # noqa
