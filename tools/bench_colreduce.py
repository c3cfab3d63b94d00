"""Dev tool: device time of the column reductions (BatchNorm statistics, bias gradient) over typical training shapes;
the scalar kernels are reached by handing in a 4-byte-misaligned view."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from curvecloudnet_b200.native import call  # noqa: E402


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


for M, C in ((400000, 128), (42144, 1024), (42144, 512), (31357, 64), (120000, 256), (8192, 256), (400000, 64), (100000, 192)):
    buf = torch.randn(M * C + 4, device="cuda")
    out = {}
    for name, off in (("v4", 0), ("scalar", 1)):
        z = buf[off:off + M * C].view(M, C)
        sums = torch.empty(2 * C, dtype=torch.float64, device="cuda")
        mean, invstd = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
        o = torch.zeros(C, device="cuda")
        t_stats = timed(lambda: call("ccn_bn_stats", z, C, M, C, 1e-5, 0.1, sums, mean, invstd, 0, 0))
        t_sum = timed(lambda: call("ccn_col_sum", z, C, M, C, o))
        dy = torch.randn(M, C, device="cuda")
        dz = torch.empty(M, C, device="cuda")
        gamma, beta = torch.ones(C, device="cuda"), torch.zeros(C, device="cuda")
        dg, db = torch.empty(C, device="cuda"), torch.empty(C, device="cuda")
        t_bwd = timed(lambda: call("ccn_bn_backward", dy, C, z, C, M, C, mean, invstd, gamma, beta, 1, 0.0, 1, sums, dz, C,
                                   dg, db, 0, 0, 0, 0, 0))
        out[name] = (t_stats, t_sum, t_bwd)
    mb = M * C * 4 / 1e6
    print("M=%7d C=%5d %7.1f MB  bn_stats v4 %6.1f us (%5.0f GB/s) scalar %6.1f us | col_sum v4 %6.1f us (%5.0f GB/s) scalar %6.1f us"
          " | bn_backward v4-reduce %6.1f us scalar-reduce %6.1f us"
          % (M, C, mb, out["v4"][0], mb / out["v4"][0] * 1e3, out["scalar"][0], out["v4"][1], mb / out["v4"][1] * 1e3,
             out["scalar"][1], out["v4"][2], out["scalar"][2]))
