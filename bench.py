"""bench.py - CurveCloudNet nuScenes-shape segmentation forward (BASELINE.json configs[3]).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one forward of the 33-step nuScenes CurveCloudNet over one synthetic 32-beam scan
(~33 k points, B=1, eval semantics of the reference's val loop).  Per rank: S distinct scans are
cycled; ranks shard scans, no data-path collective (per-scan inference has no cross-GPU dependency).

One JSON line on stdout (rank 0):
  value      points/s, whole job, inputs resident in HBM when a step starts (device-event time, max over ranks)
  e2e        points/s through the public API (ModelBase.forward) from PINNED HOST buffers, H2D of the
             scan and D2H of the logits inside the timed region
  roofline   the dominant kernel family (the shared-MLP Linear kernel): algorithmic BYTES (activation planes read
             once, weights once, outputs written once) / CUDA-event time of its launches inside the timed steps,
             against the measured HBM copy peak (every layer of this network sits below the tensor/HBM ridge);
             the tensor-pipe figures (FLOPs / time against the measured bf16 peak) are kept beside it
  cpu_baseline  the CPU oracle (port of the reference, oracle/) timed on this host on a bounded sample
--impl reference: the reference's CPU path = the oracle port (the reference itself cannot be imported
  on this image), all host threads, bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "points/sec fwd seg (nuScenes-shape CurveCloudNet)"
WORKLOAD = "nuScenes-shape CurveCloudNet segmentation forward, 32-beam synthetic scan, B=1 scan/step"


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def _ncu_traffic(key):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture (profiles/ncu_traffic.json)."""
    try:
        return float(json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[key]["dram_bytes_per_launch"])
    except Exception:
        return None


def _pin_rank_cores(local, world):
    """The per-scan path is launch-bound on the host: give every rank its own slice of the allowed cores so that the ranks'
    Python / driver threads do not migrate onto each other (BENCH_NO_PIN=1 disables)."""
    if world <= 1 or os.environ.get("BENCH_NO_PIN"):
        return None
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = len(cores) // world
        if per < 2:
            return None
        mine = cores[local * per:(local + 1) * per]
        os.sched_setaffinity(0, mine)
        return len(mine)
    except Exception:
        return None


class ClockSampler:
    def __init__(self, device_index):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(device_index), "--query-gpu=" + q,
                                       "--format=csv,noheader,nounits", "-lms", "150"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            # NVML start-up and its first query take driver locks for 0.1-0.3 s and stall CUDA calls of this process
            # (measured: e2e steps of 10 ms instead of 5 ms when that happened inside the timed passes): wait for the
            # first sample before any timed work.  The periodic queries (every 150 ms) keep running during the timed passes;
            # that first, idle sample is not counted.
            self.first = self.p.stdout.readline()
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [t.strip() for t in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "samples": len(sm),
                "reasons": sorted(reasons)}


def _cpu_model():
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.synthetic import randomize_model_
    from oracle.ccn_model import ModelBase as OracleModel
    model = build_model(OracleModel, "nuscenes")
    randomize_model_(model, 1234)
    return model.eval()


def cpu_reference_run(scans, steps, warmup, threads, budget_s=None):
    """Times the CPU oracle (port of the reference; its neighbour searches are the grid-pruned C statement
    oracle/nbr_search.c, multi-threaded) on the SAME scans, cycled in the same order with the same per-step seeds as
    the GPU arm.  A 256 MiB host buffer is written between the timed steps (the CPU-side counterpart of the L2 flush).
    -> (points/s, mean points per step, seconds per step, timed steps)."""
    torch.set_num_threads(threads)
    model = _cpu_model()
    flush = torch.empty(256 << 20, dtype=torch.uint8)
    times, npts = [], 0
    t_begin = time.perf_counter()
    with torch.no_grad():
        for i in range(-warmup, steps):
            flush.zero_()
            data = scans[(i if i >= 0 else i + warmup) % len(scans)]
            torch.manual_seed(max(i, 0))
            t0 = time.perf_counter()
            model(data)
            dt = time.perf_counter() - t0
            if i >= 0:
                times.append(dt)
                npts += data.pos.shape[0]
                if budget_s is not None and time.perf_counter() - t_begin > budget_s and len(times) >= 3:
                    break
    total = sum(times)
    return npts / total, npts / len(times), total / len(times), len(times)


def cpu_frnn_bruteforce_points_per_s(scan, K=20, radius=0.04, max_queries=4096):
    """the reference's own FRNNBruteForceCPU (third_party/FRNN/frnn/csrc/bruteforce/
    bruteforce_cpu.cpp:5-58, compiled from the reference tree into oracle/_ref/) on the scan's level-0 self-query -
    the one piece of the reference's neighbour search that exists as CPU code.  Single-threaded O(P1*P2) upstream;
    timed on a bounded number of queries against the full cloud.  None if the library was not built."""
    import ctypes
    so = os.path.join(ROOT, "oracle", "_ref", "libref_frnn_bf.so")
    if not os.path.exists(so):
        return None
    try:
        lib = ctypes.CDLL(so)
        vp = ctypes.c_void_p
        lib.ref_frnn_bruteforce_cpu.argtypes = [vp, vp, vp, vp, ctypes.c_int64, ctypes.c_int64, ctypes.c_int64, ctypes.c_int,
                                                ctypes.c_float, vp, vp]
        pos = scan.pos.contiguous()
        nq = min(max_queries, pos.shape[0])
        p1 = pos[:nq].contiguous()
        l1, l2 = torch.tensor([nq]), torch.tensor([pos.shape[0]])
        idx = torch.empty((1, nq, K), dtype=torch.int64)
        d = torch.empty((1, nq, K), dtype=torch.float32)
        t0 = time.perf_counter()
        rc = lib.ref_frnn_bruteforce_cpu(p1.data_ptr(), pos.data_ptr(), l1.data_ptr(), l2.data_ptr(), 1, nq, pos.shape[0], K,
                                         float(radius), idx.data_ptr(), d.data_ptr())
        dt = time.perf_counter() - t0
        if rc != 0:
            return None
    except Exception:
        return None
    return {"queries_per_s": nq / dt, "queries": nq, "sources": int(pos.shape[0]), "K": K, "radius": radius,
            "seconds": dt, "kind": "reference (FRNNBruteForceCPU compiled from the reference tree), 1 thread"}


def _rank0_scans(rank, n_scans, batch=1):
    from curvecloudnet_b200.synthetic import driving_scan_batch
    return [driving_scan_batch([1000 * rank + batch * s + b for b in range(batch)]) for s in range(n_scans)]


def _config(mean_pts, batch, n_scans, world):
    """The `config` object of BOTH arms (the reference arm runs the same scans in the same order)."""
    return {"workload": WORKLOAD.replace("B=1", "B=%d" % batch), "points_per_step": mean_pts, "scans_per_step": batch,
            "distinct_batches_per_rank": n_scans, "parallelism": "scan-sharded dp%d, no collective" % world,
            "l2": "256 MiB buffer written between timed steps (cache flush)"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    threads = os.cpu_count() or 1
    steps, warmup = max(args.steps, 1), max(args.warmup, 3)
    scans = _rank0_scans(0, args.scans, args.batch)
    # the whole run must end within a few minutes: past 8 minutes of CPU work the timed steps stop early (>= 3 kept)
    # and the line reports the number of steps actually timed
    pps, mean_pts, sec, done = cpu_reference_run(scans, steps, warmup, threads, budget_s=480.0)
    sample = ("oracle port (plain-torch restatement of the reference; grid-pruned C neighbour search), the same %d full scans "
              "cycled as the GPU arm, %d threads" % (len(scans), threads))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": pps, "unit": "points/s", "n_gpus": args.gpus, "steps": done,
        "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": _config(mean_pts, args.batch, args.scans, world),
        "cpu_baseline": {"value": pps, "unit": "points/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": pps, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------------------------------
# --mode train: BASELINE.json configs[1], ShapeNet-seg (grid sampling) fwd+bwd, batch 16 x 2048 points per GPU
# ------------------------------------------------------------------------------------------------------------------
TRAIN_METRIC = "points/sec fwd+bwd training step (ShapeNet-seg-shape CurveCloudNet)"
TRAIN_WORKLOAD = "ShapeNet-seg-shape CurveCloudNet training step (fwd + NLL loss + bwd + Adam), B=%d x 2048 synthetic scan-line points"
# --train-config: the other data-parallel training configurations of BASELINE.json (per-GPU batch fixed: weak scaling)
TRAIN_CONFIGS = {
    "shapenet-seg": dict(model="shapenet-seg", batch=16, lovasz=False, metric=TRAIN_METRIC, workload=TRAIN_WORKLOAD),
    "kortx": dict(model="kortx", batch=64, lovasz=False,
                  metric="points/sec fwd+bwd training step (Kortx-shape CurveCloudNet)",
                  workload="Kortx-shape CurveCloudNet training step (fwd + NLL loss + bwd + Adam), B=%d x 2048-point crops per GPU (BASELINE configs[2])"),
    "kitti": dict(model="kitti", batch=2, lovasz=True,
                  metric="points/sec fwd+bwd training step (KITTI-shape CurveCloudNet)",
                  workload="KITTI/A2D2-shape CurveCloudNet training step (fwd + NLL + Lovasz loss + bwd + Adam), B=%d x ~114 k-point 64-beam scans per GPU (BASELINE configs[4])"),
}


def _train_batch(cfg, seeds):
    from curvecloudnet_b200.synthetic import driving_scan_batch, object_scan_batch
    if cfg["model"] == "kitti":
        return driving_scan_batch(seeds, kind="kitti", n_classes=20)
    return object_scan_batch(seeds)


def cpu_train_points_per_s(batch, iters, warmup, threads):
    """The CPU oracle's training step (autograd over the restatement) on a bounded batch."""
    import torch.nn.functional as F
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.synthetic import object_scan_batch, randomize_model_
    from oracle.ccn_model import ModelBase as OracleModel
    torch.set_num_threads(threads)
    data = object_scan_batch(list(range(batch)))
    model = build_model(OracleModel, "shapenet-seg")
    randomize_model_(model, 1234)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    gen = torch.Generator().manual_seed(0)
    target = torch.randint(0, 50, (data.pos.shape[0],), generator=gen)
    times = []
    for i in range(warmup + iters):
        torch.manual_seed(i)
        t0 = time.perf_counter()
        opt.zero_grad()
        loss = F.nll_loss(F.log_softmax(model(data), 1), target)
        loss.backward()
        opt.step()
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    return data.pos.shape[0] * len(times) / sum(times), data.pos.shape[0], sum(times) / len(times)


def run_train_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    threads = os.cpu_count() or 1
    steps, warmup = max(min(args.steps, 5), 1), 1
    pps, n, sec = cpu_train_points_per_s(2, steps, warmup, threads)
    sample = "oracle port, B=2 x 2048 points per step, %d threads" % threads
    print(json.dumps({
        "impl": "reference", "metric": TRAIN_METRIC, "value": pps, "unit": "points/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": TRAIN_WORKLOAD % (args.train_batch or 16), "sample": sample},
        "cpu_baseline": {"value": pps, "unit": "points/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": pps, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}))


def run_train(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from curvecloudnet_b200 import autograd as A
    from curvecloudnet_b200 import native
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.sharding import GradReducer
    from curvecloudnet_b200.synthetic import object_scan_batch, randomize_model_
    native.load()
    tcfg = TRAIN_CONFIGS[args.train_config]
    W, K = max(args.warmup, 3), max(args.steps, 1)
    B = args.train_batch if args.train_batch > 0 else tcfg["batch"]
    WG = "ccn_linear_wgrad" if args.precision == "fp32" else "ccn_linear_wgrad_tc"
    LIN = "ccn_linear_f32" if args.precision == "fp32" else "ccn_linear_planes"
    from curvecloudnet_b200 import losses as L
    model = build_model(ModelBase, tcfg["model"], precision=args.precision)
    randomize_model_(model, 1234)
    model = model.cuda().train()
    reducer = GradReducer(model.parameters(), dist)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)
    gen = torch.Generator().manual_seed(rank)
    host, targets = [], []
    for s in range(args.scans):
        d = _train_batch(tcfg, [10000 * rank + B * s + b for b in range(B)])
        host.append(d.pin_memory())
        targets.append(torch.randint(0, model.n_out, (d.pos.shape[0],), generator=gen).pin_memory())
    dev = [(d.to("cuda"), t.to("cuda")) for d, t in zip(host, targets)]
    pts = [d.num_points for d in host]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_pass(steps, from_host, profile):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        npts = 0
        native.profile_begin((WG, LIN) if profile else None, pad_cycles=40000 if profile else 0)
        barrier()
        l0 = native.kernel_launches()
        for i in range(steps):
            flush.zero_()
            j = i % args.scans
            torch.manual_seed(i)
            ev[i][0].record()
            if from_host:
                data, tgt = host[j].to("cuda", non_blocking=True), targets[j].to("cuda", non_blocking=True)
            else:
                data, tgt = dev[j]
            reducer.prepare()
            if tcfg["lovasz"]:
                loss, _ = L.seg_loss(model(data), tgt, ignore=0, use_lovasz=True)
            else:
                loss = A.nll_loss(model(data), tgt)
            loss.backward()
            reducer.finish()
            opt.step()
            if from_host:
                loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
            ev[i][1].record()
            npts += pts[j]
        barrier()
        launches = native.kernel_launches() - l0
        prof = native.profile_end()
        return sum(a.elapsed_time(b) for a, b in ev), npts, launches, prof

    timed_pass(W, False, False)
    sampler = ClockSampler(local) if rank == 0 and not os.environ.get("BENCH_NO_CLOCKS") else None   # env: dev A/B only
    ms, npts, launches, _ = timed_pass(K, False, False)
    timed_pass(min(W, 3), True, False)                                     # warm-up of the host-fed arm
    ms_e2e, npts_e2e, _, _ = timed_pass(K, True, False)
    ms_prof, _, _, prof = timed_pass(K, False, True)
    clocks = sampler.stop() if sampler else None
    t = torch.tensor([ms, ms_e2e], dtype=torch.float64, device="cuda")
    tot = torch.tensor([npts, npts_e2e], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_max, ms_e2e_max = [float(v) for v in t.tolist()]
    tot_pts, tot_pts_e2e = [float(v) for v in tot.tolist()]
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    peaks, peak_src = _peaks()
    peak_tf = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1371.7)))
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))

    def kernel_roofline(name, kernel, traffic_key):
        # both GEMM kernels of the step sit below the tensor / HBM ridge (210 FLOP/B) at this network's widths
        # (DESIGN.md section 4): algorithmic bytes against the measured copy bandwidth, tensor-pipe view beside it
        d = prof.get(name, {"ms": 0.0, "calls": 0, "work": 0.0, "bytes": 0.0})
        sec = d["ms"] * 1e-3
        gbs = d["bytes"] / sec / 1e9 if sec > 0 else 0.0
        tf = d["work"] / sec / 1e12 if sec > 0 else 0.0
        return {"bound": "hbm", "kernel": kernel, "achieved": gbs, "peak": peak_gbs, "unit": "GB/s",
                "frac": gbs / peak_gbs if peak_gbs else None, "traffic": _ncu_traffic(traffic_key),
                "peak_source": peak_src + " (hbm_gbs)", "launches": d["calls"],
                "algorithmic_bytes_per_launch": d["bytes"] / d["calls"] if d["calls"] else None,
                "avg_launch_us": (d["ms"] * 1e3 / d["calls"]) if d["calls"] else None,
                "share_of_step": d["ms"] / ms_prof if ms_prof else None, "flops_per_step": d["work"] / K,
                "tensor_achieved_tflops": tf, "tensor_peak_tflops": peak_tf, "tensor_frac": tf / peak_tf if peak_tf else None}

    if args.precision == "fp32":
        roofline = kernel_roofline("ccn_linear_f32", "linear_f32_kernel (ccn_linear_f32)", "ccn_linear_f32:fp32")
    else:
        # the dominant kernel family of the step: forward Linear layers and data gradients (dX = dY W) run on the same kernel
        roofline = kernel_roofline(LIN, "linear_planes_kernel<%d> (ccn_linear_planes: forward + data gradients)"
                                   % (3 if args.precision == "bf16x3" else 1), "ccn_linear_planes:train:%s" % args.precision)
        roofline["wgrad"] = kernel_roofline(WG, "wgrad_tc_kernel (ccn_linear_wgrad_tc: TMA + tcgen05 MN-major + TMEM, split-M)",
                                            "ccn_linear_wgrad_tc:%s" % args.precision)
    line = {
        "metric": tcfg["metric"], "value": tot_pts / (ms_max * 1e-3), "unit": "points/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "%s Linear / curve-conv GEMMs (forward, data and weight gradients), f32 everything else" % args.precision, "data": "synthetic",
        "config": {"workload": tcfg["workload"] % B, "points_per_step": tot_pts / (K * world), "scans_per_step": B,
                   "distinct_batches_per_rank": args.scans, "optimizer": "torch.optim.Adam(fused)",
                   "parallelism": "data parallel dp%d, bucketed gradient all-reduce overlapped with backward" % world,
                   "l2": "256 MiB buffer written between timed steps (L2 flush)"},
        "e2e": {"value": tot_pts_e2e / (ms_e2e_max * 1e-3), "unit": "points/s",
                "h2d_bytes_per_step": int(sum(d.input_bytes() for d in host) / len(host) + targets[0].numel() * 8),
                "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e_max / K},
        "gpu_launches": int(launches),
        "roofline": roofline,
        "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline and args.train_config == "shapenet-seg":
        threads = os.cpu_count() or 1
        pps, n, sec = cpu_train_points_per_s(2, 2, 1, threads)
        line["cpu_baseline"] = {"value": pps, "unit": "points/s", "cores": threads, "kind": "port",
                                "sample": "oracle port, B=2 x 2048 points, 2 timed steps, %.2f s each" % sec}
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def train_variant(dist, rank, world, precision, batch=16, steps=6, warmup=4):
    """Informational line inside the default (inference) bench at EVERY N: the ShapeNet-seg training step (BASELINE.json
    configs[1], fwd + NLL + bwd + fused Adam, B = 16 x 2048 points per GPU) through sharding.GradReducer - the ONE collective
    of the design (bucketed NCCL gradient all-reduce overlapped with backward), so that the driver's scaling run sees it.
    Runs on every rank; -> dict on rank 0 (device-event time, max over ranks; points of all ranks)."""
    from curvecloudnet_b200 import autograd as A
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.sharding import GradReducer
    from curvecloudnet_b200.synthetic import object_scan_batch, randomize_model_
    model = build_model(ModelBase, "shapenet-seg", precision=precision)
    randomize_model_(model, 1234)
    model = model.cuda().train()
    reducer = GradReducer(model.parameters(), dist)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)
    data = object_scan_batch([10000 * rank + b for b in range(batch)]).to("cuda")
    tgt = torch.randint(0, 50, (data.pos.shape[0],), generator=torch.Generator().manual_seed(rank)).to("cuda")
    ev = []
    for i in range(warmup + steps):
        if i == warmup:
            torch.cuda.synchronize()
            if dist is not None:
                dist.barrier()
            torch.cuda.synchronize()
        torch.manual_seed(i)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        reducer.prepare()
        loss = A.nll_loss(model(data), tgt)
        loss.backward()
        reducer.finish()
        opt.step()
        b.record()
        if i >= warmup:
            ev.append((a, b))
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    pts = torch.tensor([float(data.pos.shape[0] * steps)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(pts, op=dist.ReduceOp.SUM)
    grad_bytes = int(reducer.flat.numel() * 4)
    reducer.remove()
    return {"workload": TRAIN_WORKLOAD % batch, "value": float(pts.item()) / (float(t.item()) * 1e-3), "unit": "points/s",
            "ms_per_step": float(t.item()) / steps, "steps": steps, "warmup": warmup, "n_gpus": world, "scaling": "weak",
            "collective": "bucketed NCCL all-reduce of %d gradient bytes per step, overlapped with backward (sharding.GradReducer)"
                          % grad_bytes if world > 1 else "none (one GPU)", "dtype": precision}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scans", type=int, default=4, help="distinct synthetic scans cycled per rank")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--batch", type=int, default=1, help="scans per step (one ModelBase.forward call)")
    ap.add_argument("--precision", default="bf16x3", choices=["fp32", "bf16x3", "bf16"],
                    help="shared-MLP Linear arithmetic: fp32 pipe | tcgen05 split-bf16 x3 (1e-4 budget) | tcgen05 bf16 (2e-2 budget)")
    ap.add_argument("--mode", default="infer", choices=["infer", "train"],
                    help="infer: nuScenes-shape forward (the headline metric); train: ShapeNet-seg fwd+bwd step (configs[1])")
    ap.add_argument("--no-batched-variant", action="store_true", help="skip the informational B=6 line (batched_variant)")
    ap.add_argument("--no-train-variant", action="store_true", help="skip the informational training-step line (train_variant)")
    ap.add_argument("--train-batch", type=int, default=0, help="scans per training step per GPU (--mode train; 0 = the config's)")
    ap.add_argument("--train-config", default="shapenet-seg", choices=["shapenet-seg", "kortx", "kitti"],
                    help="--mode train: ShapeNet-seg B=16 (configs[1]) | Kortx B=64 (configs[2]) | KITTI-shape B=2, NLL+Lovasz (configs[4])")
    args = ap.parse_args()
    if args.mode == "train":
        return run_train_reference(args) if args.impl == "reference" else run_train(args)
    if args.impl == "reference":
        return run_reference(args)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback")
    torch.cuda.set_device(local)
    _pin_rank_cores(local, world)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    from curvecloudnet_b200 import native
    from curvecloudnet_b200.configs import build_model
    from curvecloudnet_b200.model import ModelBase
    from curvecloudnet_b200.synthetic import driving_scan_batch, randomize_model_
    native.load()
    W, K = max(args.warmup, 3), max(args.steps, 1)

    # BENCH_PYTHON_EXECUTOR=1 (dev A/B): the python step executor of model.py instead of ccn_forward_run
    model = build_model(ModelBase, "nuscenes", precision=args.precision, native_executor=not os.environ.get("BENCH_PYTHON_EXECUTOR"))
    randomize_model_(model, 1234)
    model = model.cuda().eval()
    lin_name = "ccn_linear_f32" if args.precision == "fp32" else "ccn_linear_planes"
    host_scans = [s.pin_memory() for s in _rank0_scans(rank, args.scans, args.batch)]
    dev_scans = [s.to("cuda") for s in host_scans]
    pts = [s.num_points for s in host_scans]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # > 126 MB L2
    n_out = model.n_out
    pinned_out = [torch.empty((max(pts), n_out), dtype=torch.float32).pin_memory() for _ in range(2)]

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def native_exec():
        st = model.__dict__.get("_native")
        return st[1] if st else None

    fam_tot = {}

    def timed_pass(steps, from_host, profile):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        outs, npts = [], 0
        nx = native_exec()
        if profile and nx is not None:
            nx.profile(True, 20000)               # per-family CUDA events inside ccn_forward_run (csrc/executor.cu)
        native.profile_begin((lin_name,) if profile else None, pad_cycles=40000 if profile else 0)
        barrier()
        l0 = native.kernel_launches()
        with torch.no_grad():
            for i in range(steps):
                flush.zero_()                                              # L2 flush, outside the timed span
                j = i % args.scans
                torch.manual_seed(i)
                ev[i][0].record()
                data = host_scans[j].to("cuda", non_blocking=True) if from_host else dev_scans[j]
                logits = model(data)
                if profile and nx is not None:
                    for f, d in nx.profile_read().items():
                        t = fam_tot.setdefault(f, {"ms": 0.0, "calls": 0.0, "flops": 0.0, "bytes": 0.0})
                        for k in t:
                            t[k] += d[k]
                if from_host:
                    # D2H of the step's result into pinned host memory, asynchronous (double-buffered): the copy is
                    # inside the step's event pair; the host only waits for it at the next forward's first readback
                    pinned_out[i & 1][:logits.shape[0]].copy_(logits, non_blocking=True)
                ev[i][1].record()
                npts += pts[j]
        barrier()
        launches = native.kernel_launches() - l0
        prof = native.profile_end()
        if profile and nx is not None:
            nx.profile(False)
        ms = sum(a.elapsed_time(b) for a, b in ev)
        return ms, npts, launches, prof, (logits.shape if steps else None)

    timed_pass(W, False, False)                                            # warm-up
    sampler = ClockSampler(local) if rank == 0 and not os.environ.get("BENCH_NO_CLOCKS") else None   # env: dev A/B only
    ms, npts, launches, _, oshape = timed_pass(K, False, False)
    timed_pass(min(W, 3), True, False)                                     # warm-up of the host-fed arm (first H2D / D2H allocations)
    ms_e2e, npts_e2e, _, _, _ = timed_pass(K, True, False)
    # roofline leg: the same K steps once more with CUDA events around every Linear launch (the events cost host time,
    # so this pass is not the one `value` is taken from; its own step time is reported as roofline.step_ms)
    ms_prof, _, _, prof, _ = timed_pass(K, False, True)
    clocks = sampler.stop() if sampler else None

    t = torch.tensor([ms, ms_e2e], dtype=torch.float64, device="cuda")
    tot = torch.tensor([npts, npts_e2e], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_max, ms_e2e_max = [float(v) for v in t.tolist()]
    tot_pts, tot_pts_e2e = [float(v) for v in tot.tolist()]
    tv = None
    if not args.no_train_variant and args.batch == 1:
        try:
            tv = train_variant(dist, rank, world, args.precision)
        except Exception as e:                      # informational: never take the headline line down
            tv = {"error": "%s: %s" % (type(e).__name__, e)}
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peaks, peak_src = _peaks()
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    peak_tf = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1371.7)))
    KERNELS = {"linear": "linear_planes_kernel (ccn_linear_planes: per-point shared-MLP layers, TMA + tcgen05 + TMEM)",
               "edge_fused": "edge_chain_kernel (ccn_sgcnn_fused / ccn_pointconv_fused: gather -> MLP chain -> aggregate, tcgen05 + TMEM)",
               "conv": "linear_planes_kernel (ccn_conv_planes: curve convolution as one GEMM)",
               "search": "grid_build / frnn_knn / knn_exact kernels", "sample": "curve_fps / voxel_fps / fps kernels"}
    if not fam_tot:
        # python executor (fp32 precision or native_executor=False): only the Linear family is instrumented (native.profile_begin)
        lin = prof.get(lin_name, {"ms": 0.0, "calls": 0, "work": 0.0, "bytes": 0.0})
        fam_tot["linear"] = {"ms": lin["ms"], "calls": lin["calls"], "flops": lin["work"], "bytes": lin["bytes"]}
        if args.precision == "fp32":
            KERNELS["linear"] = "linear_f32_kernel (ccn_linear_f32)"

    def family_view(f):
        d = fam_tot[f]
        sec = d["ms"] * 1e-3
        tf = d["flops"] / sec / 1e12 if sec > 0 else 0.0
        gbs = d["bytes"] / sec / 1e9 if sec > 0 else 0.0
        mma = 3 if args.precision == "bf16x3" else 1
        return {"kernel": KERNELS[f], "launches": int(d["calls"]), "avg_launch_us": d["ms"] * 1e3 / d["calls"] if d["calls"] else None,
                "share_of_step": d["ms"] / ms_prof if ms_prof else None,
                "algorithmic_flops_per_step": d["flops"] / K, "algorithmic_bytes_per_step": d["bytes"] / K,
                "tensor_achieved_tflops": tf, "tensor_frac": tf / peak_tf if peak_tf else None,
                "tensor_frac_executed": mma * tf / peak_tf if peak_tf else None,
                "hbm_achieved_gbs": gbs, "hbm_frac": gbs / peak_gbs if peak_gbs else None}

    # The dominant kernel family of the step by device time.  Its bound follows from its ALGORITHMIC intensity (SURVEY 8(d):
    # FLOPs over valid edges / compulsory bytes without inter-layer intermediates) against the ridge of the measured peaks:
    # the fused edge steps and the wide per-point layers sit on the tensor side, everything else on the HBM side.
    dom = max(fam_tot, key=lambda f: fam_tot[f]["ms"])
    dv = family_view(dom)
    ridge = peak_tf * 1e12 / (peak_gbs * 1e9)
    inten = (fam_tot[dom]["flops"] / fam_tot[dom]["bytes"]) if fam_tot[dom]["bytes"] else 0.0
    tensor_bound = inten >= ridge
    roofline = {"bound": "tensor" if tensor_bound else "hbm", "kernel": dv["kernel"],
                "achieved": dv["tensor_achieved_tflops"] if tensor_bound else dv["hbm_achieved_gbs"],
                "peak": peak_tf if tensor_bound else peak_gbs, "unit": "TFLOP/s" if tensor_bound else "GB/s",
                "frac": dv["tensor_frac"] if tensor_bound else dv["hbm_frac"],
                "traffic": _ncu_traffic("%s:%s" % (dom, args.precision)),
                "peak_source": peak_src + (" (bf16_tflops_sustained)" if tensor_bound else " (hbm_gbs)"),
                "algorithmic_intensity_flop_per_byte": inten, "ridge_flop_per_byte": ridge,
                "mma_per_algorithmic_mma": 3 if args.precision == "bf16x3" else 1,
                "launches": dv["launches"], "avg_launch_us": dv["avg_launch_us"], "share_of_step": dv["share_of_step"],
                "step_ms_instrumented": ms_prof / K,
                "families": {f: family_view(f) for f in fam_tot}}
    mean_pts = tot_pts / (K * world)
    h2d = sum(s.input_bytes() for s in host_scans) / len(host_scans)
    d2h = mean_pts * (oshape[1] if oshape else 17) * 4
    line = {
        "metric": METRIC, "value": tot_pts / (ms_max * 1e-3), "unit": "points/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": {"fp32": "f32", "bf16x3": "bf16x3 (split-bf16 tensor-core MLPs, fp32-class accuracy; all else f32)",
                  "bf16": "bf16 (tensor-core MLPs, fp32 accumulate; all else f32)"}[args.precision],
        "data": "synthetic", "scans_per_s": (K * world * args.batch) / (ms_max * 1e-3),
        "config": _config(mean_pts, args.batch, args.scans, world),
        "e2e": {"value": tot_pts_e2e / (ms_e2e_max * 1e-3), "unit": "points/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e_max / K},
        "gpu_launches": int(launches), "roofline": roofline, "clocks": clocks,
        "dependent_launch": bool(native.set_dependent_launch(-1)),      # programmatic dependent launch (CCN_PDL=0: off)
    }
    if world == 1 and args.batch == 1 and not args.no_batched_variant:
        # SURVEY 8(d) config 4 also names "a batched variant B=6 (train batch)": the same scans, six per forward; reported
        # beside the headline (which stays the B=1 eval semantics), same L2 flush and event timing, 3 warm-up + 8 timed steps
        B6 = 6
        batches = [driving_scan_batch([5000 + B6 * g + b for b in range(B6)]).to("cuda") for g in range(2)]
        bpts = [int(b.pos.shape[0]) for b in batches]
        tot_ms, tot_p = 0.0, 0
        for i in range(3 + 8):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.manual_seed(i)
            e0.record()
            with torch.no_grad():
                model(batches[i % 2])
            e1.record()
            torch.cuda.synchronize()
            if i >= 3:
                tot_ms += e0.elapsed_time(e1)
                tot_p += bpts[i % 2]
        line["batched_variant"] = {"scans_per_step": B6, "value": tot_p / (tot_ms * 1e-3), "unit": "points/s",
                                   "ms_per_scan": tot_ms / (8 * B6), "steps": 8, "warmup": 3}
    if tv is not None:
        line["train_variant"] = tv
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        # bounded sample of the SAME workload: the full seed-0 scan (the first scan of the timed cycle), 1 warm-up + 3 timed
        # forwards of the oracle port on all host threads
        scan0 = _rank0_scans(0, 1)
        pps, n, sec, done = cpu_reference_run(scan0, 3, 1, threads)
        line["cpu_baseline"] = {"value": pps, "unit": "points/s", "cores": threads, "kind": "port",
                                "sample": "oracle port (grid-pruned C neighbour search), the full %d-point seed-0 scan, %d timed "
                                          "forwards, %.2f s each" % (int(n), done, sec)}
        bf = cpu_frnn_bruteforce_points_per_s(scan0[0])
        if bf is not None:
            line["cpu_baseline"]["frnn_bruteforce_cpu"] = bf
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    # stdout carries the JSON line and nothing else: NCCL writes its version banner to fd 1 from C, so fd 1 is pointed at stderr
    # for the run and the real stdout is kept aside for the line(s) main() prints.
    sys.stdout.flush()
    _real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = _real
    try:
        main()
    finally:
        _real.flush()
