"""Dev tool: forward / training-step timings of the other BASELINE.json configs (not the headline bench line).
usage: python tools/bench_configs.py  -> one JSON line per case"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from curvecloudnet_b200 import autograd as A  # noqa: E402
from curvecloudnet_b200 import losses as L  # noqa: E402
from curvecloudnet_b200.configs import build_model  # noqa: E402
from curvecloudnet_b200.model import ModelBase  # noqa: E402
from curvecloudnet_b200.synthetic import driving_scan_batch, object_scan_batch, randomize_model_  # noqa: E402


def timed(fn, warm=3, it=8):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(it)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in ev) / it


def case(name, cfg, data, train, precision="bf16x3", lovasz=False):
    torch.cuda.reset_peak_memory_stats()
    model = build_model(ModelBase, cfg, precision=precision)
    randomize_model_(model, 1234)
    model = model.cuda()
    data = data.to("cuda")
    n = data.pos.shape[0]
    if train:
        model.train()
        opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)
        tgt = torch.randint(0, model.n_out, (n,), device="cuda")

        def step():
            opt.zero_grad(set_to_none=True)
            torch.manual_seed(0)
            loss, _ = L.seg_loss(model(data), tgt, ignore=0, use_lovasz=lovasz)
            loss.backward()
            opt.step()
    else:
        model.eval()

        def step():
            with torch.no_grad():
                torch.manual_seed(0)
                model(data)
    ms = timed(step)
    print(json.dumps({"case": name, "config": cfg, "mode": "train step (fwd+loss+bwd+Adam)" if train else "forward",
                      "precision": precision, "points": n, "ms": ms, "points_per_s": n / ms * 1e3,
                      "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}), flush=True)
    del model
    torch.cuda.empty_cache()


if __name__ == "__main__":
    case("kitti-shape fwd B=1 (64 beams x 1875 az)", "kitti", driving_scan_batch([0], kind="kitti", n_classes=20), False)
    case("nuscenes fwd B=1", "nuscenes", driving_scan_batch([0]), False)
    case("shapenet-class fwd B=2x2048 (configs[0])", "shapenet-class", object_scan_batch([0, 1]), False, "fp32")
    case("kortx fwd B=64x2048 (configs[2])", "kortx", object_scan_batch(list(range(64))), False)
    case("kortx train B=64x2048 (configs[2])", "kortx", object_scan_batch(list(range(64))), True)
    case("nuscenes train B=6 NLL+Lovasz", "nuscenes", driving_scan_batch(list(range(6))), True, lovasz=True)
    case("kitti-shape train B=2 NLL+Lovasz (configs[4])", "kitti", driving_scan_batch([0, 1], kind="kitti", n_classes=20), True, lovasz=True)
