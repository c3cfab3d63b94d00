"""shim: torch_geometric (see ../README.md)"""
from . import typing  # noqa
