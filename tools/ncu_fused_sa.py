"""The fused attention PointNetConv step (ccn_pointconv_fused) at the level-1 -> level-2 `sa` shape, three launches and
nothing else: the process ncu attaches to (`-k regex:edge_chain -s 2 -c 1`).  usage: python tools/ncu_fused_sa.py [precision]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from curvecloudnet_b200 import nn, ops
from curvecloudnet_b200.nn import ForwardCache

precision = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
nn.set_precision(precision)
npl = nn.n_planes()
nq, n_src, avg, mx, dims, att_h = 5695, 17809, 11.6, 32, [128, 128, 128], 64
gen = torch.Generator().manual_seed(nq)
deg = torch.clamp((torch.rand(nq, generator=gen) * 2 * avg).long() + 1, max=mx)
row_ptr = torch.zeros(nq + 1, dtype=torch.int64)
row_ptr[1:] = torch.cumsum(deg, 0)
E = int(row_ptr[-1])
erow = torch.repeat_interleave(torch.arange(nq), deg)
ecol = torch.randint(0, n_src, (E,), generator=gen)
edges = ops.EdgeCSR(row_ptr.to(torch.int32).cuda(), erow.to(torch.int32).cuda(), ecol.to(torch.int32).cuda(), E, nq, mx)
c1, C = dims[0], dims[-1]
H = torch.randn(n_src, c1, generator=gen).cuda()
wp = torch.randn(3, c1, generator=gen).cuda()
b0 = (torch.randn(c1, generator=gen) * 0.1).cuda()
ps, pd = torch.rand(n_src, 3, generator=gen).cuda(), torch.rand(nq, 3, generator=gen).cuda()


def layers_of(d, acts, seed):
    g = torch.Generator().manual_seed(seed)
    return [((torch.randn(o, i, generator=g) / i ** 0.5).cuda(), (torch.randn(o, generator=g) * 0.1).cuda(), a, 0.01)
            for i, o, a in zip(d[:-1], d[1:], acts)]


local = layers_of(dims, [2] * (len(dims) - 2) + [0], 1)
attend = layers_of([C, att_h, C], [2, 0], 2)
chain = ops.ChainWeights(local + attend, npl)
for _ in range(3):
    y = ops.pointconv_fused(edges, chain, "attend", len(local) - 1, 2, 0.01, c1, H=H, wp=wp, bias0=b0, pos_src=ps, pos_dst=pd,
                            radius=0.02, npl=npl, cache=ForwardCache())
torch.cuda.synchronize()
print("ok", E, tuple(y.shape))
