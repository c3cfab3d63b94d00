"""ModelBase: the config-driven step executor of the reference (src/models/base.py:16-209) with the same
constructor, forward contract and state_dict keys (steps.{i}..., mlp..., lin_categorical...), running
on the sm_100a kernels.  Differences in form only: per-step tensors are kept by reference instead of
cloned (base.py:177-197 clones four tensors per step; nothing here mutates them), topology objects are
memoised per forward (nn.ForwardCache), random draws come from a `Draws` object.
"""
from __future__ import annotations

import copy

import torch

from . import modules as M
from . import native, ops
from .nn import (MLP, Draws, ForwardCache, cat_features, gemm_input, get_precision, set_precision,
                 set_weights_epoch)

_DOWN_STEPS = ("sa", "sa-geo", "sa-global", "pt-transition-down")


class Batch:
    """Minimal stand-in for torch_geometric.data.Batch (the reference's input container): attribute bag
    with x, pos, batch, curve_idxs, y[, labels]."""

    def __init__(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)

    def to(self, device, non_blocking=False):
        out = Batch()
        for k, v in self.__dict__.items():
            setattr(out, k, v.to(device, non_blocking=non_blocking) if torch.is_tensor(v) else v)
        return out

    @classmethod
    def from_data_list(cls, data_list):
        out = cls()
        n_nodes = [d.pos.shape[0] for d in data_list]
        for k in data_list[0].__dict__:
            vals = [getattr(d, k) for d in data_list]
            if vals[0] is None:
                setattr(out, k, None)
            elif torch.is_tensor(vals[0]):
                setattr(out, k, torch.stack(vals) if vals[0].dim() == 0 else torch.cat(vals, dim=0))
            else:
                setattr(out, k, vals)
        dev = data_list[0].pos.device
        out.batch = torch.cat([torch.full((n,), i, dtype=torch.int64, device=dev) for i, n in enumerate(n_nodes)])
        return out


def step_input_dims(i, name, feat_dims, in_dim, with_xyz):
    """base.py:66-84 `_get_input_dim`."""
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
        # arithmetic of the shared-MLP Linear layers: "fp32" | "bf16x3" | "bf16" (nn.set_precision); not a reference kwarg
        self.precision = kwargs.pop("precision", "fp32")
        # inference on the tensor-core precisions runs through the native step executor (csrc/executor.cu, ONE C-ABI call per
        # forward) where it covers the step list; False keeps the python executor below (same launches, same logits)
        self.native_executor = kwargs.pop("native_executor", True)
        self.mlp_func = MLP
        self.step_names = list(steps)
        self.skip_connect_state_store = list(kwargs.get("skip_connect_state_store", []))
        self.steps = torch.nn.ModuleList()
        for i, entry in enumerate(steps):
            kw = dict(kwargs)
            name = entry
            if isinstance(entry, dict):                                     # base.py:40-43
                kw.update(entry)
                name = kw.pop("step_name")
                self.step_names[i] = name
            kw["with_xyz"] = kw.get("with_xyz", False)
            dims = step_input_dims(i, name, feat_dims, in_dim, kw["with_xyz"])
            self.steps.append(self.add_step(i, name, dims, **kw))

        mk = {"dropout": 0.5, "norm": "batch_norm", "plain_last": True}     # base.py:49-64
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
        if kw.get("aggr_type") not in ("attend", "weighted-sum"):
            return None
        return MLP([c, c // 2 if halved else c, c], act="leaky_relu", bias=self.use_bias)

    def add_step(self, i, name, dims, **kw):
        """base.py:86-131."""
        b = self.use_bias
        if name == "sa":
            if kw.get("aggr_type") in ("attend", "weighted-sum") and self.version not in (1.0, 2.0):
                raise UnboundLocalError("attend_nn")       # the reference leaves it unbound (base.py:88-92)
            att = self._attend(dims[-1], kw, halved=(self.version == 2.0))
            return M.SAModule(kw["ratios"][i], kw["radii"][i], MLP(dims, bias=b), attend_nn=att, k=kw["knn"][i], **kw)
        if name == "sa-global":
            return M.GlobalSAModule(MLP(dims, bias=b), **kw)
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
        if name == "dgcnn":                                                 # base.py:98-99
            return M.DGCNNLayer(MLP(dims, bias=b), kw["knn"][i], with_xyz=kw["with_xyz"])
        if name == "dgcnn-rad":                                             # base.py:100-101
            return M.DGCNNLayerRadius(MLP(dims, bias=b), kw["radii"][i], with_xyz=kw["with_xyz"])
        raise NotImplementedError("Have not implemented step %s yet!" % name)

    # ------------------------------------------------------------------------------ forward
    def forward(self, data, draws: Draws = None, **kwargs):
        prev = get_precision()
        set_precision(self.precision)
        native.pin_stream(True)
        set_weights_epoch(self._weights_epoch())
        try:
            # `labels=` is what the run scripts pass (nuscenes_seg.py:36) and no step reads: it does not keep a forward off the
            # native executor; any other keyword (trace=, cache= ...) selects the python executor
            if self.native_executor and not self.training and self.precision != "fp32" and not (set(kwargs) - {"labels"}):
                out = self._forward_native(data, draws)
                if out is not None:
                    return out
            return self._forward(data, draws, **kwargs)
        finally:
            set_weights_epoch(None)
            native.pin_stream(False)
            set_precision(prev)

    def _apply(self, fn, *a, **k):
        self.__dict__.pop("_epoch_tensors", None)       # .to()/.cuda() replace buffer tensors
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, *a, **k):
        self.__dict__.pop("_epoch_tensors", None)
        return super().load_state_dict(*a, **k)

    def _weights_epoch(self):
        """Fingerprint of every parameter/buffer (in-place version counters + storage identity + train flag)."""
        ts = self.__dict__.get("_epoch_tensors")
        if ts is None:
            ts = list(self.parameters()) + list(self.buffers())
            self.__dict__["_epoch_tensors"] = ts
        v = 0
        for t in ts:
            v += t._version
        return (id(self), self.training, v, ts[0].data_ptr() if ts else 0, len(ts))

    def _forward_native(self, data, draws):
        """One forward through ccn_forward_run; None when the model / input is outside the native executor's scope."""
        from . import executor as X
        if isinstance(data, list):
            data = Batch.from_data_list(data)
        if not data.pos.is_cuda:
            raise TypeError("curvecloudnet_b200 runs on CUDA tensors only (there is no CPU path)")
        st = self.__dict__.get("_native")
        epoch = (self._weights_epoch(), self.precision, ops.FUSED_EDGE_STEPS, ops.CHAIN_MLP_LAYERS)
        if st is None or st[0] != epoch:
            try:
                st = (epoch, X.NativeExecutor(self))
            except X.Unsupported:
                st = (epoch, None)
            self.__dict__["_native"] = st
        if st[1] is None or getattr(data, "labels", None) is not None and hasattr(self, "lin_categorical"):
            return None
        try:
            return st[1].run(data, draws)
        except X.Unsupported:
            return None

    def _forward(self, data, draws: Draws = None, **kwargs):
        """base.py:133-175.  hist[k][i] is the INPUT of step i (the lists start with the raw input)."""
        if isinstance(data, list):
            data = Batch.from_data_list(data)
        if not data.pos.is_cuda:
            raise TypeError("curvecloudnet_b200 runs on CUDA tensors only (there is no CPU path)")
        if draws is not None:
            kwargs["draws"] = draws
        kwargs["cache"] = ForwardCache()
        trace = kwargs.pop("trace", None)     # test hook: list receiving every step's (name, x, pos, batch, curve, down_idx)
        cur = (data.x, data.pos, data.batch, data.curve_idxs)
        hist = {"x": [data.x], "pos": [data.pos], "batch": [data.batch], "curve": [data.curve_idxs]}
        down_idx, prop_stack, down_stack = [], [], []
        cats = getattr(data, "labels", None)
        batch0 = cur[2]

        for i, step in enumerate(self.steps):
            name = self.step_names[i]
            if name in ("fp", "fp-geo"):                                    # base.py:161-167,199-203
                s = down_stack.pop()
                x_skip = hist["x"][s] if hist["x"][s] is not None else hist["pos"][s]
                if name == "fp":
                    args = (cur[0], cur[1], cur[2], x_skip, hist["pos"][s], hist["batch"][s], cur[3], hist["curve"][s])
                else:
                    # base.py:166 asserts the down-sampling indices ascend; CurveFPS / sorted FPS emit them so
                    args = (cur[0], down_idx[s], x_skip, hist["pos"][s], hist["batch"][s], hist["curve"][s])
            elif name == "skip-connect":                                    # base.py:168-171,205-209
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
        if cats is not None and hasattr(self, "lin_categorical"):           # base.py:150-153
            onehot = torch.zeros((cats.shape[0], 16), dtype=torch.float32, device=feat.device)
            onehot[torch.arange(cats.shape[0], device=feat.device), cats] = 1.0   # input encoding (B x 16)
            cat_feat = self.lin_categorical(onehot)
            if self.training:
                from . import autograd as A
                feat = A.concat([feat, A.gather_rows(cat_feat, batch0)])
            else:
                feat = cat_features([feat, cat_feat], [None, batch0])
        if isinstance(self.mlp, torch.nn.Identity):
            return feat
        if self.training:
            return self.mlp(feat)
        return self.mlp(gemm_input(feat, kwargs["cache"]))
