"""Python side of the native step executor (csrc/executor.cu, `ccn_forward_run`): builds the PLAN of a ModelBase - the
step list with every Linear use-site's prepared weights (BatchNorm folded, bf16 planes, TMA descriptors) - and runs one
forward as ONE C-ABI call.  Inference, tensor-core precisions, the nuScenes / KITTI step family; anything else (training,
fp32 precision, object-scan configs, `trace=`) stays on the python executor of model.py, which the native one reproduces
bit for bit (tests/test_gpu_executor.py).  Reference: src/models/base.py:133-175 and src/models/modules/*.py.
"""
from __future__ import annotations

import ctypes

import torch

from . import modules as M
from . import native, ops
from .nn import DEFAULT_DRAWS, MLP, n_planes

MAGIC = 0xCC9E0003
ST = {"conv1d-fast-v2": 1, "sa-geo": 2, "mlp": 3, "sgcnn": 4, "skip-connect": 5, "sa": 6, "fp": 7, "fp-geo": 8}
CCN_EUNSUPPORTED, CCN_ETOPOLOGY, CCN_EVOXEL_RANGE, CCN_ENOMEM = -2, -3, -4, -5
_DRAW_FN = ctypes.CFUNCTYPE(ctypes.c_int64, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_void_p)


class Unsupported(Exception):
    """The model (or this call) is outside the native executor's scope: the python executor runs instead."""


class _PlanBuilder:
    def __init__(self, model):
        self.model = model
        self.npl = n_planes()
        if self.npl == 0:
            raise Unsupported("fp32 precision runs on the python executor")
        self.lins, self.lin_f = [], []          # int / float streams of the Linear table
        self.n_lin = 0
        self.convs, self.n_conv = [], 0
        self.chains, self.n_chain = [], 0
        self.keep = []                          # every python object whose memory the plan points into

    # ---- tables
    def linear(self, lin, widths=None, chain=-1):
        """Registers one use-site of a PreparedLinear: widths = K-segmentation of its input (None: one block).
        chain >= 0: this layer and the ones after it in its MLP run as that ccn_mlp_chain entry."""
        if widths is not None and len(widths) <= 1:
            widths = None
        tw = lin.tc(tuple(widths) if widths is not None else None)
        seg = list(widths) if widths is not None else [tw.k_true]
        if len(seg) > 4:
            raise Unsupported("more than four concatenated inputs")
        hi, lo = [0] * 4, [0] * 4
        for slot, b in enumerate((32, 64, 128, 256)):
            if b == 32 or b // 2 < tw.n:
                h, l, keep = tw.maps(b)
                hi[slot], lo[slot] = h, l
                self.keep.append(keep)
        bias = lin.b.data_ptr() if lin.b is not None else 0
        self.lins += [tw.n, tw.k_true, tw.kpad, lin.act, len(seg)] + (seg + [0] * 4)[:4] + [bias] + hi + lo + [chain]
        self.lin_f.append(float(lin.slope))
        self.keep += [lin, tw]
        self.n_lin += 1
        return self.n_lin - 1

    def tc_weight_entry(self, tw, bias, act, slope):
        """A Linear-table entry for an already prepared TcWeight (curve-conv layers)."""
        hi, lo = [0] * 4, [0] * 4
        for slot, b in enumerate((32, 64, 128, 256)):
            if b == 32 or b // 2 < tw.n:
                h, l, keep = tw.maps(b)
                hi[slot], lo[slot] = h, l
                self.keep.append(keep)
        self.lins += [tw.n, tw.k_true, tw.kpad, act, 1, tw.k_true, 0, 0, 0, bias.data_ptr() if bias is not None else 0] + hi + lo + [-1]
        self.lin_f.append(float(slope))
        self.keep += [tw, bias]
        self.n_lin += 1
        return self.n_lin - 1

    def chain(self, cw):
        if cw is None:
            return -1
        self.chains += [cw.addr, cw.n_out]
        self.keep.append(cw)
        self.n_chain += 1
        return self.n_chain - 1

    def mlp_ids(self, mlp: MLP, widths):
        layers = mlp.layers()
        self.keep.append(layers)
        # the same ccn_mlp_chain launches as MLP.forward_from (nn.MLP.chain_for)
        cw, start = mlp.chain_for(tuple(widths)) if len(layers) >= 2 else (None, 0)
        return [self.linear(l, widths if i == 0 else None, self.chain(cw) if cw is not None and i == start else -1)
                for i, l in enumerate(layers)]

    # ---- PointNetConv2 block (shared by sa / sa-geo)
    def pointconv(self, conv: M.PointNetConv2, widths):
        """widths: channel blocks of x_src (features [+ positions])."""
        if conv.aggr_type not in ops.AGG:
            raise Unsupported(conv.aggr_type)
        if conv.global_nn is not None:
            raise Unsupported("PointNetConv2 with a global_nn")
        layers = conv.local_nn.layers()
        cx = sum(widths)
        lin_x, wp = conv.decomposed(cx)
        l0 = layers[0]
        att = conv.aggr_type in ("attend", "weighted-sum")
        extra = conv.attend_nn.layers() if att else ()
        cw = conv._fused.get(layers, extra) if len(layers) >= 2 else None
        local_ids = [self.linear(l) for l in layers[1:]]
        att_ids = [self.linear(l) for l in extra]
        self.keep += [layers, extra, lin_x, wp, l0]
        iv = [self.linear(lin_x, widths) if lin_x is not None else -1, wp.data_ptr(), l0.b.data_ptr() if l0.b is not None else 0,
              l0.out_features, l0.act, ops.AGG[conv.aggr_type], self.chain(cw), len(local_ids)] + local_ids + [len(att_ids)] + att_ids
        fv = [float(l0.slope), float(conv.normalize_radius or 0.0)]
        return iv, fv

    # ---- the step list
    def build(self):
        m = self.model
        if hasattr(m, "lin_categorical") or isinstance(m.mlp, torch.nn.Identity):
            raise Unsupported("category branch / identity head")
        cx0 = m.in_dim - 3
        chan = [cx0 if cx0 > 0 else None]            # channels of hist.x[i] = input of step i (None: features are the positions)
        prop_stack, down_stack = [], []
        steps_i, steps_f, n_steps = [], [], 0
        for i, (name, step) in enumerate(zip(m.step_names, m.steps)):
            if name not in ST:
                raise Unsupported("step %s" % name)
            cur = chan[-1]
            xw = [cur] if cur else []
            iv, fv = [], []
            if name == "conv1d-fast-v2":
                layers = step._layers()
                tc_from = len(layers)
                while tc_from > 0 and step._tc_layer(layers[tc_from - 1][0], tc_from - 1 == 0):
                    tc_from -= 1
                ids = []
                for li, (w, b) in enumerate(layers):
                    taps, cin, cout = w.shape
                    tc = li >= tc_from
                    lin_id = self.tc_weight_entry(step._tc_weight(li, w), b, ops.ACT_LEAKY, 0.01) if tc else -1
                    self.convs += [taps, cin, cout, int(tc), w.data_ptr(), b.data_ptr(), lin_id]
                    self.keep += [w, b]
                    ids.append(self.n_conv)
                    self.n_conv += 1
                iv = [int(step.with_xyz), int(step.with_diff), step.kernel_size, len(layers), tc_from] + ids
                out_c = layers[-1][0].shape[2]
            elif name == "mlp":
                w = xw + ([3] if step.with_xyz or not cur else [])
                ids = self.mlp_ids(step.mlp, w)
                iv = [int(step.with_xyz), len(ids)] + ids
                out_c = step.mlp.out_channels
            elif name == "skip-connect":
                ns = step.num_skips
                take = prop_stack[-ns:]
                del prop_stack[-ns:]
                w = [cur] + [chan[t] if chan[t] else 3 for t in take]
                ids = self.mlp_ids(step.nn, w)
                iv = [ns, len(ids)] + ids
                out_c = step.nn.out_channels
            elif name == "sgcnn":
                if step.use_sparse_feat_agg or step.aggr_type != "max":
                    raise Unsupported("sparse / non-max SGCNN")
                w = xw + ([3] if step.with_xyz or not cur else [])
                lin_pq, act0, slope0 = step._decomposed()
                layers = step.nn.layers()
                cw = step._fused.get(layers, dense=True) if step.k + 1 <= 128 else None
                rest = [self.linear(l) for l in layers[1:]]
                self.keep += [layers, lin_pq]
                iv = [int(step.with_xyz), step.k, self.linear(lin_pq, w), act0, self.chain(cw), len(rest)] + rest
                fv = [float(step.r), float(slope0)]
                out_c = step.nn.out_channels
            elif name == "sa-geo":
                w = xw + ([3] if step.with_xyz or not cur else [])
                piv, pfv = self.pointconv(step.conv, w)
                iv = [int(step.with_xyz), int(bool(step.use_curve_fps))] + piv
                fv = [float(step.curve_fps_arclen or 0.0), float(step.ratio or 0.0), float(step.r)] + pfv
                out_c = step.conv.local_nn.out_channels
            elif name == "sa":
                if not step.use_fast_knn or step.downsample_type not in ("voxel", "fps"):
                    raise Unsupported("sa with %s / ball query" % step.downsample_type)
                if not cur:
                    raise Unsupported("sa on position-only features")
                piv, pfv = self.pointconv(step.conv, xw)
                iv = [0 if step.downsample_type == "voxel" else 1, step.knn] + piv
                fv = [float(step.voxel_size or 0.0), float(step.ratio or 0.0), float(step.r)] + pfv
                out_c = step.conv.local_nn.out_channels
            else:   # fp / fp-geo
                s = down_stack.pop()
                skip_c = chan[s] if chan[s] else 3
                w = [cur, skip_c] + ([3] if step.with_xyz else [])
                ids = self.mlp_ids(step.nn, w)
                iv = [step.k, int(step.with_xyz), len(ids)] + ids
                out_c = step.nn.out_channels
            store = int(name in m.skip_connect_state_store)
            steps_i += [ST[name], store, len(iv), len(fv)] + iv
            steps_f += fv
            n_steps += 1
            if store:
                prop_stack.append(i)
            if name in ("sa", "sa-geo"):
                down_stack.append(i)
            chan.append(out_c)
        out_ids = self.mlp_ids(m.mlp, [chan[-1]])
        mode = 3 if self.npl == 2 else 1
        ints = ([MAGIC, mode, cx0 if cx0 > 0 else 0, m.n_out, self.n_lin, self.n_conv, self.n_chain, n_steps] + self.lins + self.convs +
                self.chains + steps_i + [len(out_ids)] + out_ids)
        floats = self.lin_f + steps_f
        return ints, floats


class NativeExecutor:
    """One prepared model -> `ccn_forward_run`.  Holds the plan's memory (weights, descriptors) and the arena."""

    def __init__(self, model):
        self.lib = native.load()
        b = _PlanBuilder(model)
        ints, floats = b.build()
        self._keep = b.keep
        self.n_out = model.n_out
        self.in_c = max(model.in_dim - 3, 0)
        ia = (ctypes.c_int64 * len(ints))(*ints)
        fa = (ctypes.c_double * max(len(floats), 1))(*floats)
        self.lib.ccn_exec_destroy.argtypes = [ctypes.c_void_p]
        self.lib.ccn_exec_destroy.restype = None
        self.lib.ccn_exec_topology_flags.argtypes = [ctypes.c_void_p]
        self.lib.ccn_exec_topology_flags.restype = ctypes.c_int32
        h = ctypes.c_void_p()
        rc = self.lib.ccn_exec_create(ctypes.cast(ia, ctypes.c_void_p), len(ints), ctypes.cast(fa, ctypes.c_void_p), len(floats),
                                      ctypes.byref(h))
        if rc:
            raise native.CcnError("ccn_exec_create failed (code %d)" % rc)
        self.handle = h
        self.arena = None
        self._needed = ctypes.c_int64(0)

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                self.lib.ccn_exec_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    FAMILIES = ("linear", "edge_fused", "conv", "search", "sample")

    def profile(self, enable: bool, spin_cycles: int = 20000):
        """Per-family device timing of the following forwards (CUDA events around every launch of a family; the FPS
        look-ahead on the side stream is off meanwhile)."""
        rc = self.lib.ccn_exec_profile(self.handle, int(bool(enable)), int(spin_cycles))
        if rc:
            raise native.CcnError("ccn_exec_profile failed (code %d)" % rc)

    def profile_read(self):
        """{family: {"ms", "calls", "flops", "bytes"}} of the last forward (synchronises)."""
        n = len(self.FAMILIES)
        buf = (ctypes.c_double * (4 * n))()
        rc = self.lib.ccn_exec_profile_read(self.handle, ctypes.cast(buf, ctypes.c_void_p), n)
        if rc:
            raise native.CcnError("ccn_exec_profile_read failed (code %d)" % rc)
        return {f: {"ms": buf[4 * i], "calls": buf[4 * i + 1], "flops": buf[4 * i + 2], "bytes": buf[4 * i + 3]}
                for i, f in enumerate(self.FAMILIES)}

    def host_stats(self):
        """{"total_ms", "blocked_ms", "syncs"} of the last forward: wall time inside ccn_forward_run, the part of it spent waiting
        for the device at the data-dependent sizes, the number of those waits."""
        buf = (ctypes.c_double * 3)()
        self.lib.ccn_exec_host_stats(self.handle, ctypes.cast(buf, ctypes.c_void_p))
        return {"total_ms": buf[0], "blocked_ms": buf[1], "syncs": int(buf[2])}

    def _draw_fn(self, draws, errors):
        def cb(user, kind, n, out):
            try:
                if kind == 0:
                    u = draws.curve_phase()
                    ctypes.c_float.from_address(out).value = float(u.reshape(-1)[0])
                    return 0
                if kind == 1:
                    noise = draws.voxel_noise(int(n)).to(dtype=torch.float32, device="cpu").contiguous()
                    if noise.shape[0] != n:
                        raise ValueError("voxel noise draw has %d entries for %d points" % (noise.shape[0], n))
                    ctypes.memmove(out, noise.data_ptr(), 4 * int(n))
                    return 0
                return int(draws.fps_start(int(n)))
            except BaseException as e:      # exceptions must not cross the C frame
                errors.append(e)
                return 0
        return _DRAW_FN(cb)

    def run(self, data, draws=None):
        draws = draws if draws is not None else DEFAULT_DRAWS
        pos = data.pos
        n = pos.shape[0]
        if n == 0:
            raise Unsupported("empty batch")
        pos = ops._f32(pos)
        x = None
        if self.in_c:
            if data.x is None:
                raise Unsupported("model expects point features")
            x, ldx = ops._rows(data.x)
            if x.shape[1] != self.in_c:
                raise Unsupported("feature width")
        elif data.x is not None:
            raise Unsupported("model expects no point features")
        batch, curve = ops._i64(data.batch), ops._i64(data.curve_idxs)
        logits = torch.empty((n, self.n_out), dtype=torch.float32, device=pos.device)
        if self.arena is None or self.arena.device != pos.device:
            self.arena = torch.empty(max(256 << 20, n * (24 << 10)), dtype=torch.uint8, device=pos.device)
        rng = torch.get_rng_state()
        for _ in range(6):
            errors = []
            fn = self._draw_fn(draws, errors)
            try:
                native.call("ccn_forward_run", self.handle.value, x if x is not None else 0, ldx if x is not None else 0, pos, batch, curve,
                            n, self.arena, self.arena.numel(), ctypes.cast(fn, ctypes.c_void_p).value, 0, logits, self.n_out,
                            ctypes.addressof(self._needed))
            except native.CcnError as e:
                code = getattr(e, "code", None)
                if errors:
                    raise errors[0]
                if code == CCN_ENOMEM:
                    # the draws already taken are replayed: restore the generator, enlarge, run again
                    torch.set_rng_state(rng)
                    want = max(2 * self.arena.numel(), 2 * int(self._needed.value))
                    self.arena = None
                    self.arena = torch.empty(want, dtype=torch.uint8, device=pos.device)
                    continue
                if code == CCN_ETOPOLOGY:
                    ops._raise_topo(int(self.lib.ccn_exec_topology_flags(self.handle)))
                if code == CCN_EVOXEL_RANGE:
                    raise native.CcnError("voxel_fps: voxel coordinates out of the supported range (|ix| < 2^17, B <= 1024)")
                if code == CCN_EUNSUPPORTED:
                    torch.set_rng_state(rng)
                    raise Unsupported("input outside the native executor's scope")
                raise
            if errors:
                raise errors[0]
            return logits
        raise native.CcnError("ccn_forward_run: arena could not be sized")
