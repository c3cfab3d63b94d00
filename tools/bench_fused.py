"""Dev tool: device time of the fused edge-step kernels against the unfused chain they replace, at the nuScenes-shape level
sizes (CUDA events, 256 MiB L2 flush between repetitions).  usage: python tools/bench_fused.py [bf16x3|bf16] [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from curvecloudnet_b200 import nn, ops
from curvecloudnet_b200.nn import ForwardCache, PreparedLinear

precision = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
nn.set_precision(precision)
npl = nn.n_planes()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn):
    ts = []
    for i in range(reps + 2):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        if i >= 2:
            ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def layers_of(dims, acts):
    gen = torch.Generator().manual_seed(1)
    out = []
    for (k, n), a in zip(zip(dims[:-1], dims[1:]), acts):
        out.append(((torch.randn(n, k, generator=gen) / k ** 0.5).cuda(), (torch.randn(n, generator=gen) * 0.1).cuda(), a, 0.01))
    return out


print("precision", precision)
for n, K, dims in [(17809, 20, [64, 64, 64]), (5695, 20, [128, 128]), (1933, 20, [256, 256])]:
    gen = torch.Generator().manual_seed(n)
    c1 = dims[0]
    PQ = torch.randn(n, 2 * c1, generator=gen).cuda()
    idx = torch.randint(0, n, (n, K), generator=gen, dtype=torch.int32)
    cnt = torch.randint(K // 2, K + 1, (n,), generator=gen)
    idx[torch.arange(K).view(1, K) >= cnt.view(n, 1)] = -1
    idx = idx.cuda()
    layers = layers_of(dims, [1] * (len(dims) - 2) + [0])
    chain = ops.ChainWeights(layers, npl)
    lins = [PreparedLinear(w, b, a, s) for w, b, a, s in layers]

    def fused():
        return ops.sgcnn_fused(PQ, idx, 1, 0.01, chain, npl, ForwardCache())

    def unfused():
        h = ops.sgcnn_layer1(PQ, idx, 1, 0.01, npl)
        for i, lin in enumerate(lins):
            h = lin(h, "f32" if i == len(lins) - 1 else "planes")
        return ops.group_max_dense(h, idx, npl, ForwardCache())

    a, b = fused(), unfused()
    assert torch.equal(a, b)
    rows = n * (K + 1)
    flops = 2.0 * rows * sum(k * m for k, m in zip(dims[:-1], dims[1:]))
    tf, tu = timed(fused), timed(unfused)
    print("sgcnn n=%6d rows=%7d dims=%-16s fused %7.1f us (%6.1f TF/s alg)   unfused %7.1f us   x%.2f"
          % (n, rows, dims, tf, flops / tf / 1e6, tu, tu / tf))
