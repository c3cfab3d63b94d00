"""torch_geometric 2.3.0 look-alikes used by the oracle and by the import shims.

TEST INFRASTRUCTURE - see oracle/__init__.py.

MLP follows torch_geometric.nn.MLP (the `mlp_func` of src/models/base.py:32):
parameter names lins.{i}.{weight,bias} and norms.{i}.module.{weight,bias,running_mean,
running_var,num_batches_tracked}; per hidden layer Linear -> BatchNorm1d(eps 1e-5) ->
act -> dropout; the last Linear is plain when plain_last.
"""
from __future__ import annotations

import inspect
from typing import List, Optional, Sequence, Union

import torch
import torch.nn.functional as F


def _resolve_act(act):
    if act is None or isinstance(act, torch.nn.Module):
        return act
    table = {"relu": torch.nn.ReLU, "leaky_relu": torch.nn.LeakyReLU, "leakyrelu": torch.nn.LeakyReLU,
             "elu": torch.nn.ELU, "tanh": torch.nn.Tanh, "sigmoid": torch.nn.Sigmoid}
    return table[act.lower()]()


class BatchNormWrap(torch.nn.Module):
    """torch_geometric.nn.norm.BatchNorm: wraps BatchNorm1d as `.module`."""

    def __init__(self, channels, eps=1e-5, momentum=0.1):
        super().__init__()
        self.module = torch.nn.BatchNorm1d(channels, eps=eps, momentum=momentum)

    def reset_parameters(self):
        self.module.reset_parameters()

    def forward(self, x):
        return self.module(x)


class MLP(torch.nn.Module):
    def __init__(self, channel_list: Sequence[int], *, dropout: Union[float, List[float]] = 0.0,
                 act="relu", act_first: bool = False, norm: Optional[str] = "batch_norm",
                 plain_last: bool = True, bias: Union[bool, List[bool]] = True, **kwargs):
        super().__init__()
        channel_list = list(channel_list)
        assert len(channel_list) >= 2
        self.channel_list = channel_list
        self.act = _resolve_act(act)
        self.act_first = act_first
        self.plain_last = plain_last
        nl = len(channel_list) - 1
        if isinstance(dropout, float) or isinstance(dropout, int):
            dropout = [float(dropout)] * nl
            if plain_last:
                dropout[-1] = 0.0
        self.dropout = dropout
        if isinstance(bias, bool):
            bias = [bias] * nl
        self.lins = torch.nn.ModuleList(
            torch.nn.Linear(i, o, bias=b) for i, o, b in zip(channel_list[:-1], channel_list[1:], bias))
        hidden = channel_list[1:-1] if plain_last else channel_list[1:]
        self.norms = torch.nn.ModuleList(
            BatchNormWrap(h) if norm is not None else torch.nn.Identity() for h in hidden)

    @property
    def in_channels(self):
        return self.channel_list[0]

    @property
    def out_channels(self):
        return self.channel_list[-1]

    def reset_parameters(self):
        for lin in self.lins:
            lin.reset_parameters()
        for n in self.norms:
            if hasattr(n, "reset_parameters"):
                n.reset_parameters()

    def forward(self, x, *args, **kwargs):
        x = x.to(self.lins[0].weight.dtype)       # no-op in f32; lets tests run the FEATURE arithmetic in f64 (positions stay f32)
        for i, (lin, norm) in enumerate(zip(self.lins, self.norms)):
            x = lin(x)
            if self.act is not None and self.act_first:
                x = self.act(x)
            x = norm(x)
            if self.act is not None and not self.act_first:
                x = self.act(x)
            x = F.dropout(x, p=self.dropout[i], training=self.training)
        if self.plain_last:
            x = self.lins[-1](x)
            x = F.dropout(x, p=self.dropout[-1], training=self.training)
        return x


class Data:
    """Minimal torch_geometric.data.Data: attribute bag."""

    def __init__(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)

    def keys(self):
        return [k for k in self.__dict__ if not k.startswith("_")]

    def to(self, device):
        for k in self.keys():
            v = getattr(self, k)
            if torch.is_tensor(v):
                setattr(self, k, v.to(device))
        return self


class Batch(Data):
    """Minimal torch_geometric.data.Batch: concatenates per-node attributes, builds
    `batch`/`ptr`; does NOT offset curve_idxs (hence curveidx_local2global)."""

    @classmethod
    def from_data_list(cls, data_list):
        out = cls()
        keys = data_list[0].keys()
        n_nodes = [d.pos.shape[0] for d in data_list]
        for k in keys:
            vals = [getattr(d, k) for d in data_list]
            if vals[0] is None:
                setattr(out, k, None)
            elif torch.is_tensor(vals[0]):
                if vals[0].dim() == 0:
                    setattr(out, k, torch.stack(vals))
                else:
                    setattr(out, k, torch.cat(vals, dim=0))
            else:
                setattr(out, k, vals)
        out.batch = torch.cat([torch.full((n,), i, dtype=torch.int64) for i, n in enumerate(n_nodes)])
        out.ptr = torch.tensor([0] + list(torch.tensor(n_nodes).cumsum(0)), dtype=torch.int64)
        return out


class MessagePassing(torch.nn.Module):
    """Just enough of torch_geometric.nn.conv.MessagePassing.propagate for
    PointNetConv2 (point_conv.py:52) and StaticEdgeConv.forward_slow (dgcnn.py:225):
    edge_index = [source j ; target i]; *_j gathers from the source side with
    edge_index[0], *_i from the target side with edge_index[1]; aggregate() receives
    index = edge_index[1] and dim_size = number of target nodes."""

    def __init__(self, aggr="add", *, aggr_kwargs=None, flow="source_to_target", node_dim=-2,
                 decomposed_layers=1, **kwargs):
        super().__init__()
        self.aggr = aggr
        self.flow = flow
        self.node_dim = node_dim

    def propagate(self, edge_index, size=None, **kwargs):
        src, dst = edge_index[0], edge_index[1]
        n_dst = None
        margs = {}
        for name in inspect.signature(self.message).parameters:
            base, side = name[:-2], name[-2:]
            val = kwargs[base]
            if isinstance(val, (tuple, list)):
                v = val[0] if side == "_j" else val[1]
                if val[1] is not None:
                    n_dst = val[1].shape[0]
            else:
                v = val
                if val is not None:
                    n_dst = val.shape[0]
            margs[name] = None if v is None else v[src if side == "_j" else dst]
        if n_dst is None:
            for val in kwargs.values():
                if isinstance(val, (tuple, list)) and val[1] is not None:
                    n_dst = val[1].shape[0]
        out = self.message(**margs)
        return self.aggregate(out, dst, ptr=None, dim_size=n_dst)

    def aggregate(self, inputs, index, ptr=None, dim_size=None):
        """torch_geometric 2.3.0 default aggregation for a string `aggr` (DynamicEdgeConv, dgcnn.py:49, relies on it):
        torch_scatter reductions over the target index; a target without edges gets 0."""
        from . import tp_ops as tp
        if self.aggr == "max":
            return tp.scatter_max(inputs, index, 0, None, dim_size)[0]
        if self.aggr == "mean":
            return tp.scatter_mean(inputs, index, 0, None, dim_size)
        if self.aggr in ("add", "sum"):
            return tp.scatter_add(inputs, index, 0, None, dim_size)
        raise NotImplementedError("aggr %r" % (self.aggr,))
