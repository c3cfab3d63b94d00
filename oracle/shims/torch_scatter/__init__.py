"""shim: torch_scatter 2.1.1 (see ../README.md)"""
from oracle.tp_ops import scatter_add, scatter_max, scatter_min, scatter_mean  # noqa
