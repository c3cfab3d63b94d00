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
for n, K, dims in [(17809, 20, [64, 64, 64]), (5695, 20, [128, 128]), (1933, 20, [256, 256]), (580, 20, [512, 512]), (174, 20, [1024, 1024])]:
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


def random_csr(nq, n_src, avg_deg, max_deg, seed):
    gen = torch.Generator().manual_seed(seed)
    deg = torch.clamp((torch.rand(nq, generator=gen) * 2 * avg_deg).long() + 1, max=max_deg)
    row_ptr = torch.zeros(nq + 1, dtype=torch.int64)
    row_ptr[1:] = torch.cumsum(deg, 0)
    E = int(row_ptr[-1])
    erow = torch.repeat_interleave(torch.arange(nq), deg)
    ecol = torch.randint(0, n_src, (E,), generator=gen)
    return ops.EdgeCSR(row_ptr.to(torch.int32).cuda(), erow.to(torch.int32).cuda(), ecol.to(torch.int32).cuda(), E, nq, max_deg)


for name, nq, n_src, avg, mx, dims, att_h, mode in [("sa-geo", 17809, 32568, 5.2, 9, [64, 128, 192, 256], 256, "attend"),
                                                    ("sa L1->L2", 5695, 17809, 11.6, 32, [128, 128, 128], 64, "attend"),
                                                    ("sa L2->L3", 1933, 5695, 17, 32, [256, 256, 256], 128, "attend")]:
    edges = random_csr(nq, n_src, avg, mx, nq)
    gen = torch.Generator().manual_seed(7)
    c1, C = dims[0], dims[-1]
    H = torch.randn(n_src, c1, generator=gen).cuda()
    wp = torch.randn(3, c1, generator=gen).cuda()
    b0 = (torch.randn(c1, generator=gen) * 0.1).cuda()
    ps, pd = torch.rand(n_src, 3, generator=gen).cuda(), torch.rand(nq, 3, generator=gen).cuda()
    local = layers_of(dims, [2] * (len(dims) - 2) + [0])
    attend = layers_of([C, att_h, C], [2, 0])
    chain = ops.ChainWeights(local + attend, npl)
    ll = [PreparedLinear(w, b, a, s) for w, b, a, s in local]
    al = [PreparedLinear(w, b, a, s) for w, b, a, s in attend]

    def fused():
        return ops.pointconv_fused(edges, chain, mode, len(local) - 1, 2, 0.01, c1, H=H, wp=wp, bias0=b0, pos_src=ps, pos_dst=pd,
                                   radius=0.02, npl=npl, cache=ForwardCache())

    def unfused():
        h = ops.edge_layer1(H, wp, b0, ps, pd, edges, 0.02, 2, 0.01, npl)
        for i, lin in enumerate(ll):
            h = lin(h, "both" if i == len(ll) - 1 else "planes")
        msg, a_in = h
        for i, lin in enumerate(al):
            a_in = lin(a_in, "f32" if i == len(al) - 1 else "planes")
        return ops.seg_aggregate(msg, a_in, edges.row_ptr, edges.nq, mode, npl, ForwardCache())

    a, b = fused(), unfused()
    assert torch.equal(a, b)
    alld = dims + [att_h, C]
    flops = 2.0 * edges.n_edges * (sum(k * m for k, m in zip(dims[:-1], dims[1:])) + C * att_h + att_h * C)
    tf, tu = timed(fused), timed(unfused)
    print("%-10s E=%6d nq=%5d dims=%-22s fused %7.1f us (%6.1f TF/s alg)   unfused %7.1f us   x%.2f"
          % (name, edges.n_edges, nq, alld, tf, flops / tf / 1e6, tu, tu / tf))


print("per-point MLP chains (ccn_mlp_chain) against per-layer ccn_linear_planes launches")
for M, segs, dims in [(32568, [64], [64, 64, 17]), (17809, [256], [128, 128, 64]), (17809, [64, 64], [128, 128]),
                      (32568, [32, 128], [128, 64]), (32568, [64, 32, 3], [128, 128]), (17809, [64, 64], [64, 64]),
                      (5695, [128, 128], [256, 128])]:
    gen = torch.Generator().manual_seed(M)
    parts = [torch.randn(M, k, generator=gen).cuda() for k in segs]
    layers = layers_of([sum(segs)] + dims, [2] * (len(dims) - 1) + [0])
    A = ops.PlanesCat([ops.pack_planes([t], npl) for t in parts]) if len(parts) > 1 else ops.pack_planes(parts, npl)
    chain = ops.ChainWeights(layers, npl, first_seg_widths=segs)
    lins = [PreparedLinear(w, b, a, s) for w, b, a, s in layers]
    tws = [lins[0].tc(tuple(segs) if len(segs) > 1 else None)] + [l.tc() for l in lins[1:]]

    def fused():
        return ops.mlp_chain(A, chain, "both")

    def unfused():
        h = A
        for i, (lin, tw) in enumerate(zip(lins, tws)):
            h = ops.linear_planes(h, tw, lin.b, lin.act, lin.slope, "both" if i == len(lins) - 1 else "planes")
        return h

    assert torch.equal(fused()[0], unfused()[0])
    tf, tu = timed(fused), timed(unfused)
    print("mlp M=%6d in=%-14s dims=%-16s chain %6.1f us   per-layer %6.1f us   x%.2f" % (M, segs, dims, tf, tu, tu / tf))
