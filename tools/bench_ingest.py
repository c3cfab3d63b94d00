"""Scan ingest timing: raw records in host memory -> Data fields in HBM.

    python tools/bench_ingest.py [--dataset nuscenes|kitti|a2d2] [--steps 200]

Prints one JSON line: sweeps/s and points/s of `ingest_scan` from host bytes (pinned staging + H2D + ccn_ingest_scan,
training augmentation on), the device-only time of the C-ABI call (CUDA events), its algorithmic HBM bytes, and the CPU
restatement of the reference's `get()` body timed on this host for the same sweep when --cpu-arm is given (that arm
imports oracle/ccn_ingest.py: the oracle as a timed CPU baseline, like bench.py's cpu_baseline leg; off by default).
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def a2d2(args):
    """A2D2-shaped frames (5 lidars, no beam ids): ingest_a2d2_frame = ccn_a2d2_curves + ccn_a2d2_finish + ccn_augment_points."""
    from curvecloudnet_b200.ingest import A2D2Augmentation, ingest_a2d2_frame
    from curvecloudnet_b200.synthetic import a2d2_frame
    frames = [a2d2_frame(20 + s) for s in range(4)]
    np.random.seed(1)
    for i in range(8):
        ingest_a2d2_frame(frames[i % 4], frames[i % 4]["label"], A2D2Augmentation.draw(), check=False)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(args.steps):
        ingest_a2d2_frame(frames[i % 4], frames[i % 4]["label"], A2D2Augmentation.draw(), check=False)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / args.steps
    pts = float(np.mean([f["points"].shape[0] for f in frames]))
    line = {"metric": "A2D2 frame ingest (kNN-graph curve builder + normalisation + augmentation)", "frames_per_s": 1.0 / dt,
            "points_per_s": pts / dt, "ms_per_frame": dt * 1e3, "points_per_frame": pts, "steps": args.steps}
    if args.cpu_arm:
        from oracle import ccn_ingest as I
        t0 = time.perf_counter()
        for i in range(4):
            I.a2d2_ingest_frame(frames[i], I.a2d2_draw_augmentation())
        line["cpu_ms_per_frame"] = (time.perf_counter() - t0) / 4 * 1e3
        line["cpu_kind"] = "oracle port of SemA2D2.get (torch CPU + numpy, one thread of python)"
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dataset", default="nuscenes")
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--cpu-arm", action="store_true", help="also time the CPU restatement (imports oracle/)")
    args = ap.parse_args()
    if args.dataset == "a2d2":
        return a2d2(args)
    from curvecloudnet_b200 import native
    from curvecloudnet_b200.ingest import LearningMap, ScanAugmentation, ingest_scan
    from curvecloudnet_b200.synthetic import raycast_driving_scan

    nusc = args.dataset == "nuscenes"
    raws = []
    for seed in range(8):
        r = raycast_driving_scan(100 + seed) if nusc else raycast_driving_scan(100 + seed, beams=64, azimuth_steps=1875,
                                                                              elev_deg=(-24.8, 2.0))
        cols = [r.points, r.reflectance[:, None]] + ([r.beam_ids[:, None]] if nusc else [])
        rec = np.concatenate(cols, axis=1).astype(np.float32)
        lab = (np.arange(rec.shape[0]) % 32).astype(np.uint8 if nusc else np.int32)
        raws.append((rec, lab))
    mapping = {k: k % 17 for k in range(32)}
    lm = LearningMap(mapping)
    rng = np.random.RandomState(0)

    def one(i):
        rec, lab = raws[i % len(raws)]
        return ingest_scan(rec, lab, lm, args.dataset, ScanAugmentation.draw(args.dataset, rng), check=False)

    for i in range(10):
        one(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    pts = 0
    for i in range(args.steps):
        pts += one(i).pos.shape[0]
    torch.cuda.synchronize()
    sec = time.perf_counter() - t0

    native.profile_begin(("ccn_ingest_scan",), pad_cycles=20000)
    for i in range(32):
        one(i)
    prof = native.profile_end()["ccn_ingest_scan"]
    dev_us = prof["ms"] * 1e3 / prof["calls"]
    n_mean = pts / args.steps
    rec_b = 20 if nusc else 16
    # algorithmic bytes: records read once (+ labels), pos / x / y / curve ids / reorder / inverse written once
    alg = n_mean * (rec_b + (1 if nusc else 4) + 12 + 4 + 8 + 8 + 8 + 8)

    rec, lab = raws[0]
    cpu_sec = None
    if args.cpu_arm:
        from oracle import ccn_ingest as oi
        t1 = time.perf_counter()
        reps = 3
        for _ in range(reps):
            oi.ingest_scan(rec, lab, mapping, args.dataset, oi.draw_augmentation(args.dataset, rng))
        cpu_sec = (time.perf_counter() - t1) / reps
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    print(json.dumps({
        "metric": "sweeps/s ingest (raw host bytes -> Data in HBM, training augmentation)", "dataset": args.dataset,
        "value": args.steps / sec, "points_per_s": pts / sec, "points_per_sweep": n_mean, "ms_per_sweep": sec / args.steps * 1e3,
        "device_us_per_call": dev_us, "algorithmic_bytes_per_call": alg, "device_gbs": alg / dev_us / 1e3,
        "hbm_peak_gbs": peak, "cpu_port_ms_per_sweep": cpu_sec * 1e3 if cpu_sec else None,
        "cpu_port_points_per_s": rec.shape[0] / cpu_sec if cpu_sec else None,
        "speedup_vs_cpu_port": cpu_sec / (sec / args.steps) if cpu_sec else None}))


if __name__ == "__main__":
    main()
