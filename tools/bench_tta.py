"""TTA eval loop timing: 12 augmented forwards per scan, loop vs stacked.

    python tools/bench_tta.py [--precision bf16x3] [--steps 5]
Prints one JSON line per `group` (1 = the reference's loop, 4, 12): ms per scan (12 augmented copies), scans/s.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--precision", default="bf16x3")
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    from curvecloudnet_b200 import losses as L
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    model = build_model(ModelBase, "nuscenes", precision=args.precision)
    randomize_model_(model, 1234)
    model = model.cuda().eval()
    scans = [driving_scan_batch([200 + i]).to("cuda") for i in range(4)]
    for group in (1, 4, 12):
        for i in range(2):
            L.tta_forward(model, scans[i % 4], group=group)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        pts = 0
        for i in range(args.steps):
            L.tta_forward(model, scans[i % 4], group=group)
            pts += scans[i % 4].pos.shape[0]
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        print(json.dumps({"metric": "TTA eval (12 augmented forwards per scan, nuScenes shape)", "group": group,
                          "precision": args.precision, "ms_per_scan": ms, "scans_per_s": 1e3 / ms,
                          "points_per_s_x12": 12 * pts / args.steps / ms * 1e3,
                          "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}))


if __name__ == "__main__":
    main()
