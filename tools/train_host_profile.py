"""Host side of one training step (ShapeNet-seg B = 16 x 2048, bench.py --mode train's step): where the Python time goes.
Prints (1) wall / device-done time per step, (2) cProfile's top functions by own time over a few steps, (3) the number of
C-ABI calls and torch allocations per step.   Usage: python tools/train_host_profile.py [model] [batch] > profiles/..."""
import cProfile
import io
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "shapenet-seg"
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    from curvecloudnet_b200 import autograd as A
    from curvecloudnet_b200 import native
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import object_scan_batch, randomize_model_
    model = build_model(ModelBase, name, precision="bf16x3")
    randomize_model_(model, 1234)
    model = model.cuda().train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)
    gen = torch.Generator().manual_seed(5)
    batches, targets = [], []
    for s in range(2):
        d = object_scan_batch(list(range(100 + s * batch, 100 + (s + 1) * batch)))
        batches.append(d.to("cuda"))
        targets.append(torch.randint(0, model.n_out, (d.pos.shape[0],), generator=gen).cuda())

    def step(i):
        opt.zero_grad(set_to_none=True)
        loss = A.nll_loss(model(batches[i % 2]), targets[i % 2])
        loss.backward()
        opt.step()
        return loss

    for i in range(6):
        step(i)
    torch.cuda.synchronize()
    for i in range(4):
        t0 = time.perf_counter()
        step(i)
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        print("step: host return %.2f ms, device done %.2f ms" % ((t1 - t0) * 1e3, (t2 - t0) * 1e3))

    # forward / loss / backward / optimiser split of the host time
    for i in range(2):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        out = model(batches[i % 2])
        t1 = time.perf_counter()
        loss = A.nll_loss(out, targets[i % 2])
        t2 = time.perf_counter()
        loss.backward()
        t3 = time.perf_counter()
        opt.step()
        t4 = time.perf_counter()
        torch.cuda.synchronize()
        print("host split: forward %.2f  loss %.2f  backward %.2f  optimiser %.2f ms"
              % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3))

    # C-ABI calls and allocations per step
    counts = {"call": 0, "empty": 0}
    real_call, real_empty = native.call, torch.empty

    def call(fname, *a):
        counts["call"] += 1
        return real_call(fname, *a)

    def empty(*a, **k):
        counts["empty"] += 1
        return real_empty(*a, **k)

    import curvecloudnet_b200.ops as ops_mod
    mods = [m for n, m in sys.modules.items() if n.startswith("curvecloudnet_b200") and m is not None]
    saved = [(m, getattr(m, "call")) for m in mods if hasattr(m, "call")]
    for m, _ in saved:
        setattr(m, "call", call)
    torch.empty = empty
    step(0)
    torch.cuda.synchronize()
    torch.empty = real_empty
    for m, f in saved:
        setattr(m, "call", f)
    print("per step: %d C-ABI calls, %d torch.empty" % (counts["call"], counts["empty"]))
    del ops_mod

    pr = cProfile.Profile()
    pr.enable()
    for i in range(5):
        step(i)
    pr.disable()
    torch.cuda.synchronize()
    for key in ("tottime", "cumulative"):
        s = io.StringIO()
        pstats.Stats(pr, stream=s).strip_dirs().sort_stats(key).print_stats(45)
        print("==== cProfile over 5 steps, sorted by %s" % key)
        print(s.getvalue())


if __name__ == "__main__":
    main()
