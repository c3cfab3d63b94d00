"""CPU restatement of /root/reference/src/models/base.py (ModelBase).

TEST INFRASTRUCTURE - see oracle/__init__.py.  Same constructor, forward and state_dict keys
as the reference (steps.{i}..., mlp..., lin_categorical...).
"""
from __future__ import annotations

import copy

import torch
import torch.nn.functional as F

from . import ccn_modules as M
from .ccn_point_ops import Draws
from .tp_nn import MLP, Batch

_DOWN_STEPS = ("sa", "sa-geo", "sa-global", "pt-transition-down")


def step_input_dims(i, name, feat_dims, in_dim, with_xyz):
    """base.py:66-84 `_get_input_dim`: the MLP/conv channel list of step i."""
    prev = None if i == 0 else feat_dims[i - 1][-1]
    if name in ("dgcnn", "sgcnn"):
        first = [in_dim * 2] if i == 0 else [2 * (prev + 3 * with_xyz)]
    elif name in ("sa", "sa-global", "sa-geo"):
        first = [in_dim + 3 * with_xyz] if i == 0 else [prev + 3 + 3 * with_xyz]
    elif i == 0:
        first = [in_dim]
    elif name in ("skip-connect", "fp", "fp-geo"):
        first = []
    elif name in ("mlp", "conv1d-fast-v1", "conv1d-fast-v2"):
        first = [prev + 3 * with_xyz]
    else:
        raise NotImplementedError("No Module Named >> %s" % name)
    return first + list(feat_dims[i])


class ModelBase(torch.nn.Module):
    def __init__(self, in_dim, n_out, steps=(), feat_dims=(), out_mlp=dict(), **kwargs):
        super().__init__()
        self.in_dim, self.n_out = in_dim, n_out
        self.use_bias = kwargs.get("use_bias", False)
        self.version = kwargs.get("version", 2.0)
        self.mlp_func = MLP
        self.step_names = list(steps)
        self.skip_connect_state_store = list(kwargs.get("skip_connect_state_store", []))
        self.steps = torch.nn.ModuleList()
        for i, entry in enumerate(steps):
            kw = dict(kwargs)
            name = entry
            if isinstance(entry, dict):                       # base.py:40-43
                kw.update(entry)
                name = kw.pop("step_name")
                self.step_names[i] = name
            kw["with_xyz"] = kw.get("with_xyz", False)
            dims = step_input_dims(i, name, feat_dims, in_dim, kw["with_xyz"])
            self.steps.append(self._make_step(i, name, dims, **kw))

        # base.py:49-64 final MLP
        mk = {"dropout": 0.5, "norm": "batch_norm", "plain_last": True}
        om = copy.deepcopy(out_mlp)
        if isinstance(om, dict):
            hidden = om.pop("dims")
            mk.update(om)
        else:
            hidden = om
        dims = [feat_dims[-1][-1]] + list(hidden) + [n_out]
        if mk.get("with_seg_category"):
            dims[0] += 64
            self.lin_categorical = MLP([16, 64, 64])
        if mk.get("identity"):
            self.mlp = torch.nn.Identity()
        else:
            self.mlp = MLP(dims, bias=self.use_bias, **mk)

    def _attend(self, c, kw, halved):
        """attend MLPs of base.py:88-94,103-106,109-112."""
        if kw.get("aggr_type") not in ("attend", "weighted-sum"):
            return None
        mid = c // 2 if halved else c
        return MLP([c, mid, c], act="leaky_relu", bias=self.use_bias)

    def _make_step(self, i, name, dims, **kw):
        """base.py:86-131 `add_step`."""
        b = self.use_bias
        if name == "sa":
            att = self._attend(dims[-1], kw, halved=(self.version == 2.0))
            if att is not None and self.version not in (1.0, 2.0):
                raise UnboundLocalError("attend_nn")  # the reference leaves it unbound
            return M.SAModule(kw["ratios"][i], kw["radii"][i], MLP(dims, bias=b), attend_nn=att, k=kw["knn"][i], **kw)
        if name == "sa-global":
            return M.GlobalSAModule(MLP(dims, bias=b), **kw)
        if name == "dgcnn":                                                # base.py:98-99
            return M.DGCNNLayer(MLP(dims, bias=b), kw["knn"][i], with_xyz=kw["with_xyz"])
        if name == "dgcnn-rad":                                            # base.py:100-101
            return M.DGCNNLayerRadius(MLP(dims, bias=b), kw["radii"][i], with_xyz=kw["with_xyz"])
        if name == "sgcnn":
            return M.SGCNNLayer(MLP(dims, bias=b), kw["knn"][i], r=kw["radii"][i],
                                attend_nn=self._attend(dims[-1], kw, False), **kw)
        if name == "sa-geo":
            return M.CurveSAModule(kw["ratios"][i], kw["radii"][i], MLP(dims, act="leaky_relu", bias=b),
                                   attend_nn=self._attend(dims[-1], kw, False), **kw)
        if name == "conv1d-fast-v1":
            return M.SymmetricCurve1DConvFastV1(dims, kw["kernel_sizes"][i], with_xyz=kw["with_xyz"],
                                                with_diff=kw.get("with_diff", False))
        if name == "conv1d-fast-v2":
            return M.SymmetricCurve1DConvV2(dims, kw["kernel_sizes"][i], with_xyz=kw["with_xyz"],
                                            with_diff=kw.get("with_diff", False))
        if name == "skip-connect":
            return M.SkipConnect(MLP(dims, act="leaky_relu", bias=b), kw["num_skips"][i])
        if name == "fp":
            return M.FPModule(kw["knn"][i], MLP(dims, bias=b), with_xyz=kw["with_xyz"])
        if name == "fp-geo":
            return M.CurveFPModule(kw["knn"][i], MLP(dims, act="leaky_relu", bias=b), with_xyz=kw["with_xyz"])
        if name == "mlp":
            return M.SharedMLP(dims, **kw)
        raise NotImplementedError("Have not implemented step %s yet!" % name)

    # ------------------------------------------------------------------ forward
    def forward(self, data, draws: Draws = None, **kwargs):
        """base.py:133-175.  state[k][i] is the INPUT of step i (lists start with the raw input)."""
        if isinstance(data, list):
            data = Batch.from_data_list(data)
        if draws is not None:
            kwargs["draws"] = draws
        trace = kwargs.pop("trace", None)     # test hook: list receiving every step's (name, x, pos, batch, curve, down_idx)
        cur = (data.x, data.pos, data.batch, data.curve_idxs)
        hist = {"x": [data.x], "pos": [data.pos], "batch": [data.batch], "curve": [data.curve_idxs.clone()]}
        down_idx, prop_stack, down_stack = [], [], []
        if hasattr(data, "labels"):
            kwargs["shapenet-categories"] = data.labels
        batch0 = cur[2]

        for i, step in enumerate(self.steps):
            name = self.step_names[i]
            if name in ("fp", "fp-geo"):                      # base.py:161-167,199-203
                s = down_stack.pop()
                x_skip = hist["x"][s] if hist["x"][s] is not None else hist["pos"][s]
                skip = (x_skip, hist["pos"][s], hist["batch"][s], hist["curve"][s])
                if name == "fp":
                    args = (cur[0], cur[1], cur[2], skip[0], skip[1], skip[2], cur[3], skip[3])
                else:
                    d = down_idx[s]
                    assert bool(torch.all(d[1:] - d[:-1] > 0))
                    args = (cur[0], d, skip[0], skip[1], skip[2], skip[3])
            elif name == "skip-connect":                      # base.py:168-171,205-209
                n = step.num_skips
                take = prop_stack[-n:]
                del prop_stack[-n:]
                xs = [hist["x"][t] if hist["x"][t] is not None else hist["pos"][t] for t in take]
                args = ([cur[0]] + xs, cur[1], cur[2], cur[3])
            else:
                args = cur
            out = step(*args, **kwargs)
            hist["x"].append(out[0]); hist["pos"].append(out[1])
            hist["batch"].append(out[2]); hist["curve"].append(out[3])
            down_idx.append(out[5] if len(out) > 5 else None)
            if trace is not None:
                trace.append((name, out[0], out[1], out[2], out[3], down_idx[-1]))
            if name in self.skip_connect_state_store:
                prop_stack.append(i)
            if name in _DOWN_STEPS:
                down_stack.append(i)
            cur = tuple(out[:4])

        feat = cur[0]
        if "shapenet-categories" in kwargs and hasattr(self, "lin_categorical"):   # base.py:150-153
            cats = F.one_hot(kwargs["shapenet-categories"], num_classes=16).to(feat.dtype)
            feat = torch.cat([feat, self.lin_categorical(cats)[batch0]], dim=1)
        return self.mlp(feat)
