"""Dev tool: Linear-kernel microbenchmark (CUDA events, L2 flushed between launches) over the layer shapes of the
nuScenes-shape forward.  usage: python tools/bench_linear.py [bf16x3|bf16] [old|new|both] [shape-set]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from curvecloudnet_b200 import nn as cnn, ops

mode = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
which = sys.argv[2] if len(sys.argv) > 2 else "both"
SHAPES = [(373989, 64, 64), (92272, 256, 256), (92272, 128, 64), (119595, 128, 128), (40593, 256, 256), (32710, 256, 256),
          (15584, 512, 512), (12180, 512, 512), (3654, 1024, 1024), (17809, 256, 259), (32568, 64, 35), (5695, 256, 256),
          (1933, 512, 512), (580, 1024, 1024), (580, 1024, 2051), (174, 2048, 3072), (174, 2048, 1027), (32568, 17, 64)]
if len(sys.argv) > 3 and sys.argv[3] == "small":
    SHAPES = [(17809, 128, 128), (17809, 256, 259), (17809, 128, 256), (17809, 64, 128), (5695, 256, 256), (5695, 128, 256),
              (32710, 256, 256), (32710, 128, 256), (32568, 128, 128), (32568, 64, 64), (66388, 128, 128), (1933, 512, 259)]
elif len(sys.argv) > 3:
    SHAPES = SHAPES[:int(sys.argv[3])]
cnn.set_precision(mode)
npl = cnn.n_planes()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
peak_hbm = 6505.9


def timeit(fn, reps=5):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


print("%-22s %10s %10s %9s %9s %8s" % ("M,N,K", "old us", "new us", "new TF/s", "new GB/s", "hbm frac"))
for M, N, K in SHAPES:
    x = torch.randn(M, K, device="cuda")
    w = torch.randn(N, K, device="cuda") / K ** 0.5
    b = torch.randn(N, device="cuda")
    lin = cnn.PreparedLinear(w, b, act=2)
    xp = ops.pack_planes([x], npl)
    t_old = timeit(lambda: ops.linear_tc(x, lin.tc(), b, 2, 0.01, 3 if npl == 2 else 1)) if which in ("old", "both") else float("nan")
    t_new = timeit(lambda: lin(xp, "planes")) if which in ("new", "both") else float("nan")
    byts = M * (K + N) * 2 * npl + N * K * 2 * npl
    print("%-22s %10.1f %10.1f %9.1f %9.1f %8.3f" % ("%d,%d,%d" % (M, N, K), t_old, t_new, 2.0 * M * N * K / t_new / 1e6,
                                                    byts / t_new / 1e3, byts / t_new / 1e3 / peak_hbm))
