"""Dev tool: ccn_linear_planes on the big-M layer shapes of the headline forward, per N-tile width and output kind, timed
as 12 back-to-back launches between two CUDA events (launch boundaries amortised) - once on ONE input buffer (L2-hot when it
fits) and once rotating over 4 input buffers (every launch reads memory the L2 no longer holds).  Tells apart: tile-width
choice, output-format cost (fp32 / planes / both), cold vs hot operands.   usage: python tools/bench_linear2.py [bf16x3|bf16]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from curvecloudnet_b200 import nn as cnn, ops
from curvecloudnet_b200.ops import call

mode = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
cnn.set_precision(mode)
npl = cnn.n_planes()
SHAPES = [(111783, 128, 128, "f32"), (63177, 128, 128, "planes"), (63177, 128, 128, "both"), (92638, 256, 256, "planes"),
          (92638, 256, 192, "both"), (30420, 256, 256, "planes"), (40278, 256, 256, "f32"), (15251, 512, 512, "planes"),
          (3633, 1024, 1024, "planes"), (17287, 128, 128, "planes"), (32226, 128, 128, "both")]
if len(sys.argv) > 2 and sys.argv[2] == "small":        # the level 4 / 5 layers: M of one or a few row tiles, wide N and K
    SHAPES = [(173, 2048, 3072, "planes"), (173, 2048, 1024, "planes"), (173, 1024, 2048, "both"), (576, 1024, 2051, "planes"),
              (576, 1024, 1024, "planes"), (576, 512, 1024, "both"), (1918, 512, 512, "planes"), (1918, 256, 512, "both")]
REPS = 12


def run(A, tw, bias, box, out):
    M = A.shape[0]
    Y = torch.empty((M, tw.n), dtype=torch.float32, device="cuda") if out != "planes" else None
    Yp = ops.alloc_planes(M, tw.n, npl, A.buf) if out != "f32" else None
    w_hi, w_lo, _ = tw.maps(box)
    args = [A.buf, A.kpad, A.buf.stride(0), A.k] + [0, 0, 0, 0] * 3
    return lambda A=A: call("ccn_linear_planes", A.buf, A.kpad, A.buf.stride(0), A.k, *([0, 0, 0, 0] * 3), w_hi, w_lo, bias,
                            Y if Y is not None else 0, tw.n, Yp.buf if Yp is not None else 0, Yp.kpad if Yp is not None else 0,
                            Yp.buf.stride(0) if Yp is not None else 0, M, tw.n, box, 3 if npl == 2 else 1, 2, 0.01)


def timed(fns):
    for f in fns:
        f()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(REPS):
            fns[i % len(fns)]()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / REPS)
    return best


print("%-28s %4s %9s %9s   %s" % ("M,N,K,out", "box", "hot us", "rot us", "rot: TF/s alg, GB/s (in + out bytes)"))
for M, N, K, out in SHAPES:
    w = torch.randn(N, K, device="cuda") / K ** 0.5
    bias = torch.randn(N, device="cuda")
    lin = cnn.PreparedLinear(w, bias, act=2)
    tw = lin.tc()
    As = [ops.pack_planes([torch.randn(M, K, device="cuda")], npl) for _ in range(4)]
    auto = tw.pick_box_planes(M, npl)
    byts = M * K * 2 * npl + N * K * 2 * npl + M * N * ((4 if out != "planes" else 0) + (2 * npl if out != "f32" else 0))
    for box in (256, 128, 64, 32):
        if box // 2 >= N and box != 32:
            continue
        try:
            hot = timed([run(As[0], tw, bias, box, out)])
            rot = timed([run(a, tw, bias, box, out) for a in As])
        except Exception as e:                                   # a box the shape does not admit
            print("%-28s %4d   %s" % ("%d,%d,%d,%s" % (M, N, K, out), box, str(e)[:60]))
            continue
        print("%-28s %3d%s %9.1f %9.1f   %7.1f TF/s %7.0f GB/s" % ("%d,%d,%d,%s" % (M, N, K, out), box, "*" if box == auto else " ",
                                                                  hot, rot, 2.0 * M * N * K / rot / 1e6, byts / rot / 1e3))
