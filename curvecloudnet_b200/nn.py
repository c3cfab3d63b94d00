"""Host-side building blocks: the shared MLP, random draws, the per-forward cache.

`MLP` keeps torch_geometric.nn.MLP's constructor and parameter names
(lins.{i}.{weight,bias}, norms.{i}.module.{weight,bias,running_mean,running_var,num_batches_tracked})
- the `mlp_func` of the reference (src/models/base.py:32) - so reference checkpoints load with
strict=True; its forward runs our Linear kernel with BatchNorm(eval) folded into the weights and the
activation fused (ccn_linear_f32 / the tcgen05 path).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Union

import torch

from . import ops


class Draws:
    """The reference's random draws, taken from torch's GLOBAL CPU generator in the reference's call
    order: CurveFPS `torch.rand(1)` (fps_ops.py:30), VoxelFPS `torch.rand(N)` (fps_ops.py:56),
    pytorch3d's random FPS start `torch.randint(high=len, size=(1,))` per cloud (point_ops.py:64) and
    SAModule's `torch.randperm` (pointnet2.py:51).  Seeding torch identically therefore reproduces the
    reference's sampling; tests may substitute recorded values."""

    def curve_phase(self) -> torch.Tensor:
        return torch.rand(1)

    def voxel_noise(self, n: int) -> torch.Tensor:
        return torch.rand(n)

    def fps_start(self, length: int) -> int:
        return int(torch.randint(high=int(length), size=(1,)).item())

    def randperm(self, n: int) -> torch.Tensor:
        return torch.randperm(n)


DEFAULT_DRAWS = Draws()


class ForwardCache:
    """Per-forward memo of the topology objects the reference recomputes in every step
    (curveidx_local2global / batch2ptr, point_ops.py:20-54): keyed by the identity of the id tensors."""

    def __init__(self):
        self._sptr = {}
        self._topo = {}
        self.grids = {}     # (pos2 ptr, n2, radius) -> hashed grid          (ops.frnn_knn)
        self.knn = {}       # (pos1 ptr, pos2 ptr, n1, n2, K, radius) -> idx  (ops.frnn_knn)
        self.planes = {}    # (ptr, rows, cols, ld) -> activation planes of an fp32 matrix

    def planes_for(self, t: torch.Tensor, npl: int) -> "ops.Planes":
        """Activation planes of an fp32 feature/position matrix: registered by the kernel that produced it
        (put_planes) or packed here, once per forward."""
        key = (t.data_ptr(), t.shape[0], t.shape[1], t.stride(0))
        hit = self.planes.get(key)
        if hit is None:
            hit = (ops.pack_planes([t], npl), t)             # keep the tensor alive: data_ptr stays unique
            self.planes[key] = hit
        return hit[0]

    def put_planes(self, t: torch.Tensor, pl: "ops.Planes"):
        self.planes[(t.data_ptr(), t.shape[0], t.shape[1], t.stride(0))] = (pl, t)

    def sample_ptr(self, batch: torch.Tensor) -> ops.SamplePtr:
        key = (batch.data_ptr(), batch.shape[0])
        hit = self._sptr.get(key)
        if hit is None:
            hit = (ops.sample_ptr(batch), batch)      # keep the tensor alive: data_ptr stays unique
            self._sptr[key] = hit
        return hit[0]

    def put_sample_ptr(self, batch: torch.Tensor, sptr: ops.SamplePtr):
        self._sptr[(batch.data_ptr(), batch.shape[0])] = (sptr, batch)

    def topology(self, batch: torch.Tensor, curve: torch.Tensor) -> ops.CurveTopology:
        key = (batch.data_ptr(), curve.data_ptr(), batch.shape[0])
        hit = self._topo.get(key)
        if hit is None:
            topo = ops.curve_csr(batch, curve)
            hit = (topo, batch, curve)
            self._topo[key] = hit
            self.put_sample_ptr(batch, topo.sptr)
        return hit[0]


def get_cache(kwargs) -> ForwardCache:
    c = kwargs.get("cache")
    return c if c is not None else ForwardCache()


_ACTS = {"relu": (ops.ACT_RELU, 0.0), "leaky_relu": (ops.ACT_LEAKY, 0.01), "leakyrelu": (ops.ACT_LEAKY, 0.01)}


class BatchNorm(torch.nn.Module):
    """torch_geometric.nn.norm.BatchNorm: BatchNorm1d held as `.module`."""

    def __init__(self, channels: int, eps: float = 1e-5, momentum: float = 0.1):
        super().__init__()
        self.module = torch.nn.BatchNorm1d(channels, eps=eps, momentum=momentum)

    def reset_parameters(self):
        self.module.reset_parameters()


def fold_batchnorm(weight: torch.Tensor, bias: Optional[torch.Tensor], bn: Optional[torch.nn.BatchNorm1d]):
    """(W, b) of `BatchNorm_eval(x W^T + b)` as one affine map: W' = s W, b' = (b - mean) s + beta,
    s = gamma / sqrt(var + eps).  Weight preparation, not part of the per-scan path."""
    w = weight.detach().float()
    b = bias.detach().float() if bias is not None else torch.zeros(w.shape[0], device=w.device)
    if bn is not None:
        s = bn.weight.detach() / torch.sqrt(bn.running_var.detach() + bn.eps)
        w = w * s.view(-1, *([1] * (w.dim() - 1)))
        b = (b - bn.running_mean.detach()) * s + bn.bias.detach()
    return w.contiguous(), b.contiguous()


# Set by ModelBase.forward for the duration of a forward: one model-wide fingerprint of every parameter / buffer
# version, so that the per-MLP "did my weights change" check is a single comparison on the per-scan path.
_EPOCH = [None]


def set_weights_epoch(epoch):
    _EPOCH[0] = epoch


_PRECISION = {"mode": "fp32"}
PRECISIONS = ("fp32", "bf16x3", "bf16")


def set_precision(mode: str):
    """Arithmetic of the shared-MLP Linear layers:
       "fp32"    FP32-pipe kernel (ccn_linear_f32), the 1e-4 reference path;
       "bf16x3"  tcgen05 tensor cores, split-bf16 x3 (fp32-class accuracy, 1e-4 budget);
       "bf16"    tcgen05 tensor cores, plain bf16 operands, fp32 accumulation (2e-2 budget)."""
    if mode not in PRECISIONS:
        raise ValueError("precision must be one of %s" % (PRECISIONS,))
    _PRECISION["mode"] = mode


def get_precision() -> str:
    return _PRECISION["mode"]


def n_planes() -> int:
    """Activation planes of the current precision mode (0 = fp32 mode: no planes)."""
    return {"fp32": 0, "bf16x3": 2, "bf16": 1}[_PRECISION["mode"]]


def cat_features(parts, idxs=None, cache=None):
    """torch.cat(parts, dim=1) (block k row-gathered by idxs[k]) as the input of an MLP.
    Tensor-core modes: with a ForwardCache every block becomes its own K-segment of the Linear kernel (the planes of
    a tensor are made once per forward - by its producer when that is one of our kernels - and the concatenation
    is never materialised); without a cache one pack kernel writes the concatenated planes.  fp32 mode: an fp32 matrix."""
    npl = n_planes()
    if npl:
        if cache is not None and idxs is None and len(parts) <= 4:
            return ops.PlanesCat([cache.planes_for(p, npl) for p in parts])
        return ops.pack_planes(parts, npl, idxs)
    if idxs is not None:
        parts = [p if ix is None else ops.gather_rows(p, ix) for p, ix in zip(parts, idxs)]
    return ops.concat_cols(list(parts)) if len(parts) > 1 else parts[0]


def gemm_input(x, cache=None):
    """An fp32 feature matrix as the A operand of a Linear: its (cached) activation planes on the tensor-core modes."""
    npl = n_planes()
    if npl and cache is not None and torch.is_tensor(x):
        return cache.planes_for(x, npl)
    return x


def mlp_to_features(mlp, inp, cache=None):
    """Run an MLP whose result is a step output (consumed by later concatenations / Linears): on the tensor-core
    modes the last layer writes the fp32 matrix AND its activation planes, which are registered in the cache."""
    if n_planes() and cache is not None:
        y, yp = mlp(inp, out="both")
        cache.put_planes(y, yp)
        return y
    return mlp(inp)


class PreparedLinear:
    """One Linear layer ready for the kernels: fp32 (W, b) with BatchNorm folded, activation, and (lazily)
    the bf16 hi/lo copies + TMA descriptors of the tensor-core path.  Weight preparation only."""

    __slots__ = ("w", "b", "act", "slope", "_tc")

    def __init__(self, w, b, act=ops.ACT_NONE, slope=0.01):
        self.w, self.b, self.act, self.slope, self._tc = w.contiguous(), b, act, slope, None

    @property
    def out_features(self):
        return self.w.shape[0]

    @property
    def in_features(self):
        return self.w.shape[1]

    def tc(self, widths=None):
        """bf16 weight copies + TMA descriptors, per K-segmentation of the input (None = one block)."""
        if self._tc is None:
            self._tc = {}
        tw = self._tc.get(widths)
        if tw is None:
            tw = self._tc[widths] = ops.prepare_tc_weight(self.w, widths)
        return tw

    def __call__(self, x, out: str = "f32"):
        """x: fp32 (M, K) tensor or ops.Planes.  out: "f32" | "planes" | "both" (planes only on the tensor-core modes)."""
        mode = _PRECISION["mode"]
        if mode == "fp32":
            assert out == "f32" and torch.is_tensor(x)
            return ops.linear(x, self.w, self.b, self.act, self.slope)
        if x.__class__ is ops.PlanesCat:
            if len(x.segs) > 1:
                return ops.linear_planes(x, self.tc(x.widths), self.b, self.act, self.slope, out)
            x = x.segs[0]
        elif torch.is_tensor(x):
            x = ops.pack_planes([x], n_planes())
        return ops.linear_planes(x, self.tc(), self.b, self.act, self.slope, out)


class MLP(torch.nn.Module):
    def __init__(self, channel_list: Sequence[int], *, dropout: Union[float, List[float]] = 0.0, act="relu",
                 act_first: bool = False, norm: Optional[str] = "batch_norm", plain_last: bool = True,
                 bias: Union[bool, List[bool]] = True, **kwargs):
        super().__init__()
        channel_list = list(channel_list)
        assert len(channel_list) >= 2
        if act_first:
            raise NotImplementedError("act_first MLPs do not occur on the CurveCloudNet path")
        if norm not in (None, "batch_norm", "BatchNorm"):
            raise NotImplementedError("norm=%r" % (norm,))
        if isinstance(act, str):
            if act.lower() not in _ACTS:
                raise NotImplementedError("activation %r" % act)
            self.act_name = act.lower()
        else:
            raise NotImplementedError("activation must be 'relu' or 'leaky_relu'")
        self.channel_list = channel_list
        self.plain_last = plain_last
        nl = len(channel_list) - 1
        if isinstance(dropout, (float, int)):
            dropout = [float(dropout)] * nl
            if plain_last:
                dropout[-1] = 0.0
        self.dropout = dropout
        if isinstance(bias, bool):
            bias = [bias] * nl
        self.lins = torch.nn.ModuleList(
            torch.nn.Linear(i, o, bias=b) for i, o, b in zip(channel_list[:-1], channel_list[1:], bias))
        hidden = channel_list[1:-1] if plain_last else channel_list[1:]
        self.norms = torch.nn.ModuleList(BatchNorm(h) if norm is not None else torch.nn.Identity() for h in hidden)
        self._folded = None
        self._folded_key = None
        self._folded_epoch = None

    @property
    def in_channels(self):
        return self.channel_list[0]

    @property
    def out_channels(self):
        return self.channel_list[-1]

    @property
    def num_layers(self):
        return len(self.lins)

    def reset_parameters(self):
        for lin in self.lins:
            lin.reset_parameters()
        for n in self.norms:
            if hasattr(n, "reset_parameters"):
                n.reset_parameters()

    # ------------------------------------------------------------------ folded parameters
    def _key(self):
        k = [self.training]
        for i, lin in enumerate(self.lins):
            k.append(lin.weight._version)
            k.append(lin.weight.data_ptr())
            if i < len(self.norms) and isinstance(self.norms[i], BatchNorm):
                m = self.norms[i].module
                k.append(m.weight._version + m.running_mean._version + m.running_var._version + m.bias._version)
        return tuple(k)

    def layers(self):
        """[PreparedLinear] with BatchNorm(eval) folded; cached until a parameter changes."""
        if self.training:
            raise RuntimeError("MLP.layers() is the folded inference form; training goes through autograd.mlp_train")
        epoch = _EPOCH[0]
        if epoch is not None and self._folded is not None and self._folded_epoch == epoch:
            return self._folded
        key = self._key()
        self._folded_epoch = epoch
        if self._folded is None or key != self._folded_key:
            out = []
            act, slope = _ACTS[self.act_name]
            for i, lin in enumerate(self.lins):
                bn = None
                if i < len(self.norms) and isinstance(self.norms[i], BatchNorm):
                    bn = self.norms[i].module
                w, b = fold_batchnorm(lin.weight, lin.bias, bn)
                hidden = i < len(self.lins) - 1 or not self.plain_last
                out.append(PreparedLinear(w, b, act if hidden else ops.ACT_NONE, slope))
            self._folded, self._folded_key = out, key
        return self._folded

    def forward(self, x, *args, out: str = "f32", **kwargs):
        if self.training:
            from . import autograd as A
            return A.mlp_train(self, x)
        return self.forward_from(x, 0, out)

    def forward_from(self, x, first: int, out: str = "f32"):
        """Run layers first.. (the first layer(s) having been applied in a fused/decomposed form).  On the tensor-core
        modes the activations between layers are bf16 planes; the last layer writes `out`."""
        layers = self.layers()[first:]
        if not layers:
            if out != "f32" or not torch.is_tensor(x):
                raise NotImplementedError("one-layer edge MLP on the planes path (no shipped config has one)")
            return x
        tc = _PRECISION["mode"] != "fp32"
        inner = "planes" if tc else "f32"
        if tc and first == 0 and len(layers) >= 2:
            # per-point MLP: the profitable suffix of its layers as ONE launch (ops.mlp_chain_plan), same bits either way
            if torch.is_tensor(x):
                x = ops.pack_planes([x], n_planes())
            elif x.__class__ is ops.PlanesCat and len(x.segs) == 1:
                x = x.segs[0]
            widths = x.widths if x.__class__ is ops.PlanesCat else (x.k,)
            cw, start = self.chain_for(widths)
            if cw is not None:
                for lin in layers[:start]:
                    x = lin(x, inner)
                return ops.mlp_chain(x, cw, out)
        for lin in layers[:-1]:
            x = lin(x, inner)
        return layers[-1](x, out)

    def chain_for(self, widths):
        """(ops.ChainWeights, start layer) of this MLP for an input of K-blocks `widths`, or (None, 0): weight preparation,
        cached per folded-weight version and precision."""
        layers = self.layers()
        npl = n_planes()
        memo = getattr(self, "_chains", None)
        if memo is None or memo[0] is not layers:
            memo = self._chains = (layers, {})
        key = (tuple(widths), npl, ops.CHAIN_MLP_LAYERS)
        hit = memo[1].get(key)
        if hit is None:
            plan = ops.mlp_chain_plan(layers, widths) if npl else None
            if plan is None:
                hit = (None, 0)
            else:
                start, segw = plan
                hit = (ops.ChainWeights([(l.w, l.b, l.act, l.slope) for l in layers[start:]], npl, first_seg_widths=segw), start)
            memo[1][key] = hit
        return hit
