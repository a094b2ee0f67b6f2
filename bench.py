#!/usr/bin/env python
"""Benchmark of the hot path: RPN Proposal+NMS images/s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one proposal-op pass over a batch of 8 synthetic 600x1000 images per GPU
(BASELINE.json configs[1]: 38x63 map, 9 anchors, pre-NMS 6000, NMS 0.7, post-NMS 300).
N > 1 is launched by torchrun, one rank per GPU; images shard across ranks with no
collective in the timed path ("scaling": "weak").  Rank 0 prints ONE JSON line.

  value        images/s, inputs resident in HBM, CUDA events on the launching stream,
               barrier + synchronize on both sides, max over ranks.
  e2e          images/s through the C-ABI host entry points (mxr_proposal_submit_host /
               _wait_host): pinned host inputs in the reference's layout, H2D + kernel + D2H
               inside the timed region.
  roofline     the fused kernel's algorithmic bytes over ms_per_step against MEASURED_PEAKS.json;
               roofline_alu: IoU pairs counted by an un-timed instrumented launch x 15 ops over
               the measured NMS-phase time against the measured ALU peak.
  configs      the other BASELINE.json workloads (anchor target, proposal target, WIDER stress,
               the reference's TRAIN / alternate-training proposal settings, test-time NMS),
               each with its time, algorithmic bytes and roofline fraction.
  cpu_baseline the reference's proposal path on the host cores (all cores and one core): the real
               reference through oracle/ref_shim.py where its tree is mounted ("reference"),
               else oracle/restate.py, its numpy restatement ("port").
  --impl reference  times only that CPU path, with the same workload string / warm-up policy.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIG = "voc_test"
BATCH = 8
METRIC = "Proposal+NMS images/sec @600x1000"
WORKLOAD = ("voc_test: 38x63 map x 9 anchors (21546), pre-NMS 6000, NMS 0.7, post-NMS 300, batch 8 images per GPU per "
            "step (BASELINE.json configs[1])")
PRECONDITION_MS = 60.0  # un-timed GPU spin-up before the warm-up steps (see run_gpu_arm)
CHECK_IMAGES = 64  # global images of the world-size-independent result check (8 GPUs x one batch)


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def warmup_steps(args):
    return max(args.warmup, 3)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_model():
    try:
        with open("/proc/cpuinfo") as f:
            for ln in f:
                if ln.lower().startswith("model name"):
                    return ln.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


# -----------------------------------------------------------------------------------------
# CPU arm: the reference's proposal path on the host cores
# -----------------------------------------------------------------------------------------
def reference_kind():
    """"reference": the real reference code is importable here (oracle/ref_shim.py: $MXRCNN_REF, default
    /root/reference -- never the case on the GPU box); "port": oracle/restate.py."""
    try:
        from oracle import ref_shim
        return "reference" if ref_shim.available() else "port"
    except Exception:
        return "port"


def _cpu_worker(args):
    seeds, cfg_name, kind = args
    from oracle import restate, synth
    cfg = synth.CONFIGS[cfg_name]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    data = [synth.proposal_inputs(s, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"]) for s in seeds]
    if kind == "reference":
        from oracle import make_golden, ref_shim
        R = ref_shim.ref()
        t0 = time.perf_counter()
        for s, (cls, box, info) in zip(seeds, data):
            make_golden.run_reference_proposal(R, cls, box, info, cfg, False, s)  # ProposalOperator.forward itself
        return time.perf_counter() - t0
    t0 = time.perf_counter()
    for cls, box, info in data:
        restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"],
                                cfg["post"], cfg["thresh"], cfg["min_size"])
    return time.perf_counter() - t0


def cpu_throughput(n_images, procs, kind, warm_images=0, seed0=900000):
    """images/s of the CPU proposal path with `procs` worker processes, each looping over its share of the images
    (per-image independence is the reference's only parallelism, rcnn/loader.py:270-281)."""
    import multiprocessing as mp
    for v in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
        os.environ[v] = "1"
    seeds = [seed0 + i for i in range(n_images)]
    if procs <= 1:
        _cpu_worker(([seed0 - 1], CONFIG, kind))
        t0 = time.perf_counter()
        _cpu_worker((seeds, CONFIG, kind))
        return n_images / (time.perf_counter() - t0)
    shares = [s for s in (seeds[i::procs] for i in range(procs)) if s]
    ctx = mp.get_context("fork")
    with ctx.Pool(len(shares)) as pool:
        warm = max(1, warm_images // len(shares))
        pool.map(_cpu_worker, [([seed0 - 1 - j for j in range(warm)], CONFIG, kind)] * len(shares))  # imports, page-in
        t0 = time.perf_counter()
        pool.map(_cpu_worker, [(s, CONFIG, kind) for s in shares])
        wall = time.perf_counter() - t0
    return n_images / wall


def cpu_baseline_block(n_images_all, kind):
    cores = host_cores()
    allc = cpu_throughput(n_images_all, cores, kind)
    one_n = max(4, min(24, n_images_all // max(cores, 1)))
    one = cpu_throughput(one_n, 1, kind)
    src = ("the reference's own ProposalOperator.forward (rcnn/rpn/proposal.py:39-149) through oracle/ref_shim.py"
           if kind == "reference" else "oracle/restate.py (numpy restatement of rcnn/rpn/proposal.py + helper/processing/nms.py)")
    return {"value": allc, "unit": "images/s", "cores": cores, "kind": kind,
            "sample": "%d images of the same workload, %s, one process per core" % (n_images_all, src),
            "one_core": {"value": one, "unit": "images/s", "cores": 1, "sample": "%d images, one process" % one_n},
            "cpu_model": cpu_model(), "numpy": np.__version__,
            "note": "excludes MXNet's asnumpy / assign copies around the callback (favours the reference)"}


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    kind = reference_kind()
    cores = host_cores()
    steps = max(1, min(args.steps, 400))  # bounded: ~0.1 s of CPU work per image and core
    warm = warmup_steps(args)
    n = steps * BATCH
    t0 = time.perf_counter()
    v = cpu_throughput(n, cores, kind, warm_images=warm * BATCH, seed0=800000)
    wall = n / v
    src = "the reference's ProposalOperator.forward via oracle/ref_shim.py" if kind == "reference" else \
        "oracle/restate.py (numpy restatement of rcnn/rpn/proposal.py + helper/processing/nms.py)"
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": 1e3 * wall / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_per_gpu": BATCH},
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": cores, "kind": kind, "cpu_model": cpu_model(),
                         "numpy": np.__version__,
                         "sample": "%d steps x %d images, %s, the images of all steps spread over one worker process per "
                                   "core" % (steps, BATCH, src)},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": time.perf_counter() - t0,
    }
    emit(line)


# -----------------------------------------------------------------------------------------
# clocks
# -----------------------------------------------------------------------------------------
class ClockSampler(object):
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.first = index, [], None, 0

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append(ln.strip())

    def mark(self):
        self.first = len(self.rows)

    def wait_first_row(self, timeout_s):
        t0 = time.perf_counter()
        while self.proc and not self.rows and time.perf_counter() - t0 < timeout_s:
            time.sleep(0.01)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows[self.first:]:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# -----------------------------------------------------------------------------------------
# GPU arm
# -----------------------------------------------------------------------------------------
def proposal_bytes(cfg, B):
    """mandatory I/O per image = fg scores 4N + deltas 16N + im_info 12 + rois 20 post + scores 4 post."""
    A = len(cfg["scales"]) * len(cfg["ratios"])
    N = A * cfg["H"] * cfg["W"]
    return B * (20 * N + 12 + 24 * cfg["post"])


def ncu_metrics():
    """Per-launch counters of the committed `ncu --set full` capture of the same command (profiles/r02_ncu_metrics.json,
    written by tools/ncu_extract.py from the raw page): never measured under this run."""
    p = os.path.join(ROOT, "profiles", "r02_ncu_metrics.json")
    if os.path.exists(p):
        try:
            with open(p) as f:
                return json.load(f)
        except Exception:
            return None
    return None


class ProposalRunner(object):
    """Device-resident proposal calls on one stream for one workload."""

    def __init__(self, lib, _lib, handle, dev, cfg, B, cls_d, box_d, info_d, want_keep=True):
        import torch
        from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
        self.lib, self._lib, self.h, self.cfg, self.B = lib, _lib, handle, cfg, B
        A = len(cfg["scales"]) * len(cfg["ratios"])
        self.A, self.post = A, cfg["post"]
        self.anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
        self.anc = self.anchors.ctypes.data_as(_lib.c_double_p)
        self.p = _lib.ProposalParams(B, A, cfg["H"], cfg["W"], 0, 0, float(cfg["feat_stride"]), cfg["pre"], self.post,
                                     cfg["thresh"], float(cfg["min_size"]), 0)
        self.cls_d, self.box_d, self.info_d = cls_d, box_d, info_d  # (pool, B, ...) tensors
        self.pool = cls_d.shape[0]
        self.rois = torch.zeros((B * self.post, 5), dtype=torch.float32, device=dev)
        self.scores = torch.zeros((B * self.post,), dtype=torch.float32, device=dev)
        self.keep = torch.zeros((B, self.post), dtype=torch.int32, device=dev) if want_keep else None
        self.nk = torch.zeros((B, 2), dtype=torch.int32, device=dev)
        self.stream = torch.cuda.Stream(dev)
        self.sp = int(self.stream.cuda_stream)
        # everything the launch loop passes, resolved once: the loop itself is one ctypes call per step (tensor indexing
        # and data_ptr() cost more host time per step than the kernel takes on the GPU)
        self.in_ptrs = [(cls_d[j].data_ptr(), box_d[j].data_ptr(), info_d[j].data_ptr()) for j in range(self.pool)]
        self.out_ptrs = (self.rois.data_ptr(), self.scores.data_ptr(),
                         self.keep.data_ptr() if self.keep is not None else None, None, self.nk.data_ptr())
        self.p_ref = ctypes.byref(self.p)
        self.fn = lib.mxr_proposal_forward

    def step(self, i):
        a, b, c = self.in_ptrs[i % self.pool]
        o = self.out_ptrs
        st = self.fn(self.h.ptr, self.sp, self.p_ref, a, b, c, self.anc, self.post, o[0], o[1], o[2], o[3], o[4])
        if st != 0:
            self._lib.check(st)

    def timed(self, first, n):
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        for i in range(n):
            self.step(first + i)
        e1.record(self.stream)
        self.stream.synchronize()
        return e0.elapsed_time(e1) / n  # ms per call

    def instrumented(self, i=0):
        """Un-timed launches: IoU pairs evaluated / settled in float64 per launch, and the kernel's own phase stamps."""
        import torch
        self.h.count_pairs(True)
        self.step(i)
        self.h.count_pairs(False)
        pr = self.h.pairs_read(self.B).astype(np.int64)
        self.h.profile_enable(True)
        acc = []
        for r in range(4):
            self.step(i + r)
            acc.append(self.h.proposal_stamps(self.B).astype(np.int64))
        self.h.profile_read()
        self.h.profile_enable(False)
        torch.cuda.synchronize()
        s = np.mean(np.stack(acc[1:]), axis=0) if len(acc) > 1 else acc[0]
        m = s.mean(axis=0)
        ph = {"decode_filter": (m[2] - m[1]) * 1e-3, "bin_tables": (m[3] - m[2]) * 1e-3,
              "materialise_first_window": (m[6] - m[3]) * 1e-3, "nms_stripes": (m[7] - m[6]) * 1e-3,
              "tail": (m[8] - m[7]) * 1e-3, "stripes_walked": float(m[9]), "windows": float(m[14]),
              "first_stripe": {"stage": (m[10] - m[6]) * 1e-3, "bit_blocks": (m[11] - m[10]) * 1e-3,
                               "barrier": (m[12] - m[11]) * 1e-3, "resolve": (m[13] - m[12]) * 1e-3,
                               "fixed_point_rounds": float(m[15])}}
        return {"pairs_per_launch": int(pr[:, 0].sum()), "float64_pairs_per_launch": int(pr[:, 1].sum())}, ph


def secondary_configs(lib, _lib, dev, peak_gbs, alu):
    """The other BASELINE.json workloads: device-resident proposal variants, and the target / NMS helpers
    through their Python mirrors (host arrays in and out; the kernels' own time comes from the library's per-launch
    CUDA events)."""
    import torch
    from oracle import synth
    out = {}
    for name in ("voc_train", "voc_alt", "wider"):
        cfg = synth.CONFIGS[name]
        B = BATCH
        c, b_, inf = synth.proposal_batch(list(range(500, 500 + B)), cfg)
        h = _lib.Handle(dev.index or 0)
        r = ProposalRunner(lib, _lib, h, dev, cfg, B, torch.from_numpy(c).to(dev)[None], torch.from_numpy(b_).to(dev)[None],
                           torch.from_numpy(inf).to(dev)[None], want_keep=False)
        for i in range(3):
            r.step(i)
        torch.cuda.synchronize()
        ms = r.timed(0, 20)
        cnt, ph = r.instrumented()
        alg = proposal_bytes(cfg, B)
        gbs = alg / (ms * 1e-3) / 1e9
        gops = cnt["pairs_per_launch"] * 15 / (ph["nms_stripes"] * 1e-6) / 1e9 if ph["nms_stripes"] > 0 else None
        A = len(cfg["scales"]) * len(cfg["ratios"])
        out["proposal_" + name] = {
            "workload": "proposal %s: N=%d, pre %d, post %d, batch %d per call, device-resident" %
                        (name, A * cfg["H"] * cfg["W"], cfg["pre"], cfg["post"], B),
            "us_per_call": ms * 1e3, "images_per_s": B / (ms * 1e-3), "algorithmic_bytes_per_call": alg,
            "hbm": {"achieved_gbs": gbs, "frac": gbs / peak_gbs},
            "alu": {"pairs_evaluated_per_call": cnt["pairs_per_launch"], "float64_pairs_per_call": cnt["float64_pairs_per_launch"],
                    "ops_per_pair": 15, "achieved_gops": gops, "frac_of_f32_peak": gops / alu["f32"] if gops else None,
                    "nms_phase_us": ph["nms_stripes"]},
            "stripes_walked": ph["stripes_walked"], "nkeep_min": int(r.nk[:, 0].min())}
        h.close()
        del r
    # ---- helpers through the Python mirrors: wall clock with host arrays + kernel time from the library's events ----
    from mxrcnn_b200.helper.processing.nms import nms, nms_batched
    from mxrcnn_b200.rcnn.minibatch import assign_anchor
    from mxrcnn_b200.rcnn.rpn.proposal_target import _sample_rois
    from mxrcnn_b200.rcnn.tester import detections
    dh = _lib.default_handle()

    def measure(fn, n=30):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        wall_us = (time.perf_counter() - t0) / n * 1e6
        dh.profile_enable(True)
        for _ in range(10):
            fn()
        prof = dh.profile_read()
        dh.profile_enable(False)
        k_us = sum(t for t, _ in prof.values()) / 10 * 1e3
        launches = sum(c for _, c in prof.values()) / 10.0
        return wall_us, k_us, launches

    def row(workload, fn, alg_bytes, ref):
        wall_us, k_us, launches = measure(fn)
        gbs = alg_bytes / (k_us * 1e-6) / 1e9 if (alg_bytes and k_us > 0) else None
        return {"workload": workload, "us_per_call_host_arrays": wall_us, "kernel_us_per_call": k_us,
                "kernel_launches_per_call": launches, "algorithmic_bytes_per_call": alg_bytes,
                "hbm": {"achieved_gbs": gbs, "frac": gbs / peak_gbs if gbs else None}, "reference": ref}

    info = np.array([[600, 1000, 1.6]], np.float32)
    gt64 = synth.gt_boxes(4, G=64)
    out["anchor_target"] = row("configs[2]: assign_anchor 38x63x9 anchors (8151 inside) x 64 gt, 256 sampled, fg 0.5",
                               lambda: assign_anchor((1, 18, 38, 63), gt64, info), 1121672, "rcnn/minibatch.py:204-395")
    from mxrcnn_b200.rcnn.loader import anchor_labels
    gts8 = [synth.gt_boxes(40 + i, G=g_) for i, g_ in enumerate((64, 6, 1, 40, 12, 3, 25, 7))]
    infos8 = [info] * 8
    out["anchor_target_batch8"] = row("loader hand-off: anchor_labels for 8 images (1..64 gt) into device batch tensors, "
                                      "two launches per batch", lambda: anchor_labels((1, 18, 38, 63), gts8, infos8, device=dev),
                                      8 * 1121672, "rcnn/loader.py:293-329 + rcnn/minibatch.py:204-395")
    out["anchor_target_batch8"]["us_per_image"] = out["anchor_target_batch8"]["us_per_call_host_arrays"] / 8.0
    gt8 = synth.gt_boxes(5, G=8)
    allr = np.vstack((synth.jittered_rois(5, gt8, R=2000), np.hstack((np.zeros((8, 1), np.float32), gt8[:, :4]))))
    out["proposal_target"] = row("configs[3]: _sample_rois 2008 rois x 8 gt, 128 rois, fg 0.25, 21 classes",
                                 lambda: _sample_rois(allr, gt8, 32, 128, 21, "TRAIN"), 172256,
                                 "rcnn/rpn/proposal_target.py:135-196")
    per_class = [synth.class_dets(900 + c, n=300) for c in range(20)]
    out["per_class_nms_20x300"] = row("test-time per-class NMS 0.3, 20 classes x 300 dets f64, one batched launch",
                                      lambda: nms_batched(per_class, 0.3), 20 * 300 * 40 + 20 * 300 * 4, "rcnn/tester.py:50-56")
    d2000 = synth.class_dets(77, n=2000)
    out["nms_2000"] = row("configs[4]: nms 0.3 over 2000 dets f64", lambda: nms(d2000, 0.3), 2000 * 40 + 2000 * 4,
                          "helper/processing/nms.py:4-38")
    from mxrcnn_b200.helper.dataset.imdb import class_overlaps, recall_cover
    from mxrcnn_b200.helper.processing.nms import nest
    d300 = synth.class_dets(78, n=300)
    out["nest_300"] = row("demo post-processing: nest(dets, 0.9) over 300 dets f64 (SURVEY 8f N2)", lambda: nest(d300, 0.9),
                          300 * 40 + 300 * 4, "helper/processing/nms.py:40-70")
    gt12 = synth.gt_boxes(6, G=12)
    props = synth.jittered_rois(6, gt12, R=2000)[:, 1:5].astype(np.float64)
    gcls = np.arange(12) % 20 + 1
    out["roidb_overlaps_2000x12"] = row("offline roidb: best gt / IoU per proposal, 2000 proposals x 12 gt (SURVEY 8f N3)",
                                        lambda: class_overlaps(props, gt12[:, :4], gcls, 21), 2000 * 32 + 12 * 32 + 2000 * 12,
                                        "helper/dataset/imdb.py:45-55")
    out["recall_cover_2000x12"] = row("evaluate_recall matching loop, 2000 proposals x 12 gt (SURVEY 8f N3)",
                                      lambda: recall_cover(props, gt12[:, :4].astype(np.float64)), 2000 * 32 + 12 * 32 + 12 * 8,
                                      "helper/dataset/imdb.py:153-170")
    r, d, sc, shp, scl = synth.box_head_inputs(7)
    out["box_head_300x21"] = row("test-time box head: 300 rois x 21 classes (bbox_pred, clip, threshold, per-class NMS, cap)",
                                 lambda: detections(sc, r, d, shp, scl), 300 * 21 * 4 + 300 * 16 + 300 * 80 * 4,
                                 "rcnn/detector.py:73-74, rcnn/tester.py:50-65")
    return out


def run_gpu_arm(args, rank, world, local_rank):
    import torch
    from oracle import synth
    from mxrcnn_b200 import _lib, shard

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        # stdout carries the ONE JSON line: whatever NCCL logs goes to a per-process file instead
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/mxrcnn_b200_nccl_%h_%p.log")
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=dev)
        # barriers of the timed regions go through a host-side (gloo) group: an NCCL barrier is a GPU kernel, and a
        # foreign kernel between the warm-up steps and the first timed step evicts this kernel's code from the
        # instruction caches; the ranks are synchronised all the same
        host_group = dist.new_group(backend="gloo")
    cfg = synth.CONFIGS[CONFIG]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    B = BATCH
    post = cfg["post"]
    warm = warmup_steps(args)
    lib = _lib.load()
    h = _lib.Handle(local_rank)
    # ---- synthetic input pool: larger than L2 (126 MB), rotated so no step re-reads cached inputs.  Batch i of rank r
    #      holds the global images (i * world + r) * B ... + B - 1; an image's content depends on its global index only ----
    per_batch = B * (2 * A + 4 * A) * cfg["H"] * cfg["W"] * 4
    pool = max(4, int(np.ceil(160e6 / per_batch)))
    log("[bench] rank %d: building %d input batches (%.0f MB > L2)" % (rank, pool, pool * per_batch / 1e6))
    cls_h = torch.empty((pool, B, 2 * A, cfg["H"], cfg["W"]), dtype=torch.float32).pin_memory()
    box_h = torch.empty((pool, B, 4 * A, cfg["H"], cfg["W"]), dtype=torch.float32).pin_memory()
    info_h = torch.empty((pool, B, 3), dtype=torch.float32).pin_memory()

    def global_seed(g):
        return 1000003 + g

    for i in range(pool):
        seeds = [global_seed((i * world + rank) * B + j) for j in range(B)]
        c, b_, inf = synth.proposal_batch(seeds, cfg)
        cls_h[i].copy_(torch.from_numpy(c))
        box_h[i].copy_(torch.from_numpy(b_))
        info_h[i].copy_(torch.from_numpy(inf))
    cls_d, box_d, info_d = cls_h.to(dev), box_h.to(dev), info_h.to(dev)
    run = ProposalRunner(lib, _lib, h, dev, cfg, B, cls_d, box_d, info_d)
    stream, sp, p, anc = run.stream, run.sp, run.p, run.anc
    rois, scores, keep, nk = run.rois, run.scores, run.keep, run.nk
    torch.cuda.synchronize(dev)
    step = run.step

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier(group=host_group)
        torch.cuda.synchronize(dev)

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    lib.mxr_reserve_proposal(h.ptr, ctypes.byref(p))
    # The clock sampler (nvidia-smi -lms 100) is started here and samples through every timed region below.  Its process
    # start-up and NVML initialisation take ~0.2 s of driver calls; a K = 20 timed region is 0.5 ms, so the sampler is
    # given until its first row to settle instead of initialising on top of the timed launches.
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        sampler.wait_first_row(3.0)
        sampler.mark()  # rows before this point were taken on an idle GPU: not part of the report
    # Pre-conditioning (un-timed, before the W warm-up steps): a fresh GPU idles at ~1.1 GHz and a timed region of
    # K = 20 steps lasts half a millisecond, so the same step is run for PRECONDITION_MS first to bring the clocks up.
    t_pre = time.perf_counter()
    n_pre = 0
    while (time.perf_counter() - t_pre) * 1e3 < PRECONDITION_MS:
        for i in range(50):
            step(n_pre + i)
        n_pre += 50
        torch.cuda.synchronize(dev)
    assert int(nk[:, 0].min()) == post and int(nk[:, 1].max()) == 0, nk
    launches_per_step = h.last_launch_count()
    # the W warm-up steps come last before the timed region: nothing else (no torch kernel of a check, no copy) runs
    # between them and the first timed step
    for i in range(warm):
        step(i)

    # ---- timed region: device-resident inputs ----
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for i in range(args.steps):
        step(warm + i)
    e1.record(stream)
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    value = world * B * args.steps / (ms * 1e-3)

    # ---- same K steps with two batches in flight (not the headline): a second handle (own workspace, own stream, own
    #      outputs) runs the odd steps next to the even ones ----
    h_b = _lib.Handle(local_rank)
    lib.mxr_reserve_proposal(h_b.ptr, ctypes.byref(p))
    run_b = ProposalRunner(lib, _lib, h_b, dev, cfg, B, cls_d, box_d, info_d)

    def step_pair(i):
        (run if i % 2 == 0 else run_b).step(i)

    for i in range(4):
        step_pair(i)
    barrier()
    f0, f1, fb = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    f0.record(stream)
    run_b.stream.wait_event(f0)
    for i in range(args.steps):
        step_pair(warm + i)
    fb.record(run_b.stream)
    stream.wait_event(fb)
    f1.record(stream)
    barrier()
    ms_pair = max_over_ranks(f0.elapsed_time(f1))
    assert int(run_b.nk[:, 0].min()) == post
    two_in_flight = {"value": world * B * args.steps / (ms_pair * 1e-3), "unit": "images/s",
                     "ms_per_step": ms_pair / args.steps,
                     "how": "steps alternate between two handles / streams; same kernel, same K steps"}
    h_b.close()
    del run_b

    # ---- e2e: host buffers through the C-ABI host entry points ----
    cls_np, box_np, info_np = cls_h.numpy(), box_h.numpy(), info_h.numpy()
    o_rois = np.empty((B * post, 5), np.float32)
    o_scores = np.empty((B * post,), np.float32)
    o_keep = np.empty((B, post), np.int32)
    o_nk = np.zeros((B, 2), np.int32)

    host_ptrs = [(_lib.hptr(cls_np[j]), _lib.hptr(box_np[j]), _lib.hptr(info_np[j])) for j in range(pool)]
    o_ptrs = (_lib.hptr(o_rois), _lib.hptr(o_scores), _lib.hptr(o_keep), _lib.hptr(o_nk))
    p_ref = ctypes.byref(p)

    def step_host(i):
        a_, b__, c_ = host_ptrs[i % pool]
        st = lib.mxr_proposal_forward_host(h.ptr, p_ref, a_, b__, c_, anc, post, o_ptrs[0], o_ptrs[1], o_ptrs[2], None,
                                           o_ptrs[3])
        if st != 0:
            _lib.check(st)

    # pipelined form of the same entry point: two slots, so the H2D copy of batch i+1 overlaps the kernel of batch i;
    # every step still copies its own inputs in and its own results out inside the timed region
    outs = [(np.empty((B * post, 5), np.float32), np.empty((B * post,), np.float32), np.empty((B, post), np.int32),
             np.zeros((B, 2), np.int32)) for _ in range(2)]
    n_fg, n_box = B * A * cfg["H"] * cfg["W"], B * 4 * A * cfg["H"] * cfg["W"]
    # caller-packed variant of the inputs (fg scores | deltas | im_info in one block): a smaller pool is enough
    ppool = min(pool, 8)
    pack_h = torch.empty((ppool, n_fg + n_box + B * 3), dtype=torch.float32).pin_memory()
    for i in range(ppool):
        pack_h[i, :n_fg] = cls_h[i][:, A:].reshape(-1)
        pack_h[i, n_fg:n_fg + n_box] = box_h[i].reshape(-1)
        pack_h[i, n_fg + n_box:] = info_h[i].reshape(-1)
    pack_np = pack_h.numpy()
    pack_ptrs = [_lib.hptr(pack_np[j]) for j in range(ppool)]
    outs_ptrs = [tuple(_lib.hptr(x) for x in o) for o in outs]

    def submit(i, packed=False):
        if packed:
            st = lib.mxr_proposal_submit_host_packed(h.ptr, i % 2, p_ref, pack_ptrs[i % ppool], anc, post)
        else:
            a_, b__, c_ = host_ptrs[i % pool]
            st = lib.mxr_proposal_submit_host(h.ptr, i % 2, p_ref, a_, b__, c_, anc, post)
        if st != 0:
            _lib.check(st)

    def wait(i):
        q = outs_ptrs[i % 2]
        st = lib.mxr_proposal_wait_host(h.ptr, i % 2, q[0], q[1], q[2], None, q[3])
        if st != 0:
            _lib.check(st)
        assert outs[i % 2][3][0, 0] == post

    def run_pipelined(first, count, packed=False):
        submit(first, packed)
        for i in range(first + 1, first + count):
            submit(i, packed)
            wait(i - 1)
        wait(first + count - 1)

    for i in range(warm):
        step_host(i)
    run_pipelined(0, max(warm, 24))  # un-timed: pinned blocks touched, both slots' streams and copy engines warm
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step_host(warm + i)
    torch.cuda.synchronize(dev)
    e2e_sync_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    t0 = time.perf_counter()
    run_pipelined(warm, args.steps)
    torch.cuda.synchronize(dev)
    e2e_s_rank = time.perf_counter() - t0
    e2e_s = max_over_ranks(e2e_s_rank)
    barrier()
    # both forms processed the same last batch: identical keep lists
    assert np.array_equal(outs[(warm + args.steps - 1) % 2][2], o_keep)
    run_pipelined(0, 4, packed=True)
    barrier()
    t0 = time.perf_counter()
    run_pipelined(warm, args.steps, packed=True)
    torch.cuda.synchronize(dev)
    e2e_packed_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_sync_value = world * B * args.steps / e2e_sync_s
    clocks = sampler.stop() if rank == 0 else None
    e2e_value = world * B * args.steps / e2e_s
    N = A * cfg["H"] * cfg["W"]
    h2d = B * (4 * N + 16 * N + 12)  # fg half of cls_prob + deltas + im_info (the bg half is never copied)
    d2h = B * post * (20 + 4 + 4 + 4) + B * 8

    # what the host link gives plain pinned -> device copies, all ranks at once (the e2e bound on this box): one step's
    # input bytes per copy, and 64 MiB per copy (the link's ceiling)
    def plain_copy_gbs(nbytes, reps):
        pin = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        dst = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        with torch.cuda.stream(stream):
            for _ in range(3):
                dst.copy_(pin, non_blocking=True)
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            c0.record(stream)
            for _ in range(reps):
                dst.copy_(pin, non_blocking=True)
            c1.record(stream)
        stream.synchronize()
        v = nbytes * reps / (c0.elapsed_time(c1) * 1e-3) / 1e9
        barrier()
        return v

    link_step = plain_copy_gbs(h2d, 50)
    link_big = plain_copy_gbs(64 << 20, 8)
    my_h2d_gbs = h2d * args.steps / e2e_s_rank / 1e9
    per_rank = [(my_h2d_gbs, link_big, link_step)]
    if dist is not None:
        t = torch.tensor([my_h2d_gbs, link_big, link_step], dtype=torch.float64, device=dev)
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t)
        per_rank = [tuple(float(x) for x in q.cpu()) for q in parts]

    # ---- per-kernel profile (separate pass, CUDA events around every launch on the launching stream) ----
    prof_steps = min(args.steps, 50)
    h.profile_enable(True)
    for i in range(prof_steps):
        step(i)
    prof = h.profile_read()
    h.profile_enable(False)
    counts, phases = run.instrumented()

    # ---- measured ALU peaks (in-repo micro-benchmark, no FMA contraction) for the NMS roofline ----
    alu = {}
    for nm, code in (("f64", 1), ("f32", 0)):
        gv = ctypes.c_double(0)
        _lib.check(lib.mxr_bench_alu_peak(h.ptr, code, ctypes.byref(gv)))
        alu[nm] = gv.value

    # ---- saturating batch (not the headline): 256 images in one launch = 132 MB of inputs (> L2): the throughput
    #      asymptote of the same kernel; and one image per launch: the latency floor ----
    sat = lat1 = None
    if rank == 0:
        SB = 256
        reps = (SB + pool * B - 1) // (pool * B)
        cls_s = cls_d.reshape(pool * B, 2 * A, cfg["H"], cfg["W"]).repeat(reps, 1, 1, 1)[:SB].contiguous()[None]
        box_s = box_d.reshape(pool * B, 4 * A, cfg["H"], cfg["W"]).repeat(reps, 1, 1, 1)[:SB].contiguous()[None]
        info_s = info_d.reshape(pool * B, 3).repeat(reps, 1)[:SB].contiguous()[None]
        rs = ProposalRunner(lib, _lib, h, dev, cfg, SB, cls_s, box_s, info_s, want_keep=False)
        for i in range(3):
            rs.step(i)
        torch.cuda.synchronize(dev)
        sat_ms = rs.timed(0, 20)
        assert int(rs.nk[:, 0].min()) == post
        sat = {"batch": SB, "value": SB / (sat_ms * 1e-3), "unit": "images/s", "ms_per_step": sat_ms,
               "op_hbm_gbs": proposal_bytes(cfg, SB) / (sat_ms * 1e-3) / 1e9}
        sat["op_hbm_frac"] = sat["op_hbm_gbs"] / measured_peaks()[0]
        del rs, cls_s, box_s, info_s
        r1 = ProposalRunner(lib, _lib, h, dev, cfg, 1, cls_d[:, :1].contiguous(), box_d[:, :1].contiguous(),
                            info_d[:, :1].contiguous(), want_keep=False)
        for i in range(3):
            r1.step(i)
        torch.cuda.synchronize(dev)
        l_ms = r1.timed(0, 50)
        lat1 = {"batch": 1, "us_per_call": l_ms * 1e3, "value": 1.0 / (l_ms * 1e-3), "unit": "images/s",
                "note": "mandatory I/O of one image is 0.07 us at HBM peak: a single image is a latency chain"}
        del r1

    # ---- parity spot check of one batch against the oracle (not timed) ----
    ok = None
    if rank == 0:
        from oracle import restate
        step(0)
        torch.cuda.synchronize(dev)
        c0_, b0_, i0_ = cls_np[0][:1], box_np[0][:1], info_np[0][:1]
        st_ = restate.proposal_stages(c0_, b0_, i0_, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"],
                                      post, cfg["thresh"], cfg["min_size"], nms_stop_after=post)
        ok = bool(np.array_equal(keep[0].cpu().numpy(), st_["keep"]))

    # ---- world-size independence (harness only, not timed): the same CHECK_IMAGES global images at every N, sharded
    #      image g -> rank g % world (rcnn/loader.py:270-281), results all-gathered (NCCL) and checksummed in global order ----
    mine = shard.shard_indices(CHECK_IMAGES, rank, world)
    loc = {"rois": [], "scores": [], "keep": [], "nkeep": []}
    for b0 in range(0, len(mine), B):
        ids = mine[b0:b0 + B]
        ids_p = ids + [ids[-1]] * (B - len(ids))
        c, b_, inf = synth.proposal_batch([global_seed(7000000 + g) for g in ids_p], cfg)
        chk = ProposalRunner(lib, _lib, h, dev, cfg, B, torch.from_numpy(c).to(dev)[None], torch.from_numpy(b_).to(dev)[None],
                             torch.from_numpy(inf).to(dev)[None])
        chk.step(0)
        torch.cuda.synchronize(dev)
        n = len(ids)
        loc["rois"].append(chk.rois.reshape(B, post, 5)[:n, :, 1:].clone())
        loc["scores"].append(chk.scores.reshape(B, post)[:n].clone())
        loc["keep"].append(chk.keep[:n].clone())
        loc["nkeep"].append(chk.nk[:n, :1].clone())
        del chk
    gathered = {k_: shard.all_gather_outputs(torch.cat(v), CHECK_IMAGES) for k_, v in loc.items()}
    checksum = None
    if rank == 0:
        parts = [gathered["keep"].cpu().numpy().astype(np.int64), gathered["nkeep"].cpu().numpy().astype(np.int64),
                 gathered["rois"].cpu().numpy().view(np.int32).astype(np.int64).reshape(CHECK_IMAGES, -1),
                 gathered["scores"].cpu().numpy().view(np.int32).astype(np.int64)]
        checksum = {"value": "%016x" % shard.checksum(np.concatenate([x.reshape(CHECK_IMAGES, -1) for x in parts], axis=1)),
                    "images": CHECK_IMAGES,
                    "over": "keep lists, nkeep, rois bits and score bits of %d global images in global order; equal at "
                            "every --gpus N" % CHECK_IMAGES,
                    "gathered_with": "nccl all_gather" if dist is not None else "single rank"}

    if rank == 0:
        peak, peak_src = measured_peaks()
        alg = proposal_bytes(cfg, B)
        per_kernel = {}
        step_ms = 0.0
        for name, (tot, cnt) in prof.items():
            per_kernel[name] = {"ms_per_step": tot / prof_steps, "launches_per_step": cnt / prof_steps,
                                "us_per_launch_event_bracketed": 1e3 * tot / cnt}
            step_ms += tot / prof_steps
        for name in per_kernel:
            per_kernel[name]["share"] = per_kernel[name]["ms_per_step"] / step_ms if step_ms else None
        dom = max(per_kernel, key=lambda k_: per_kernel[k_]["ms_per_step"]) if per_kernel else None
        ms_step = ms / args.steps
        ncu = ncu_metrics()
        traffic = None
        if ncu and ncu.get("workload") == WORKLOAD:
            traffic = ncu.get("dram_bytes_read", 0) + ncu.get("dram_bytes_write", 0)
        ach = alg / (ms_step * 1e-3) / 1e9
        roof = {"kernel": "proposal_fused_kernel (mxrcnn_b200/csrc/proposal_fused.cuh)", "bound": "hbm", "achieved": ach,
                "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic,
                "traffic_source": "profiles/r02_ncu_metrics.json (ncu --set full of the same command, per launch)" if traffic else None,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": alg, "us_per_launch": ms_step * 1e3,
                "timed_as": "ms_per_step (one launch per step, back to back on one stream)",
                "share_of_step": per_kernel[dom]["share"] if dom else None,
                "note": "0.5 us of HBM time per batch against a chain of dependent phases: the op is latency-bound at this "
                        "batch by construction (SURVEY 8d); saturating_batch shows the asymptote"}
        # ALU view of the NMS phase (BASELINE.json: "NMS by SM FP32/ALU-pipe utilisation"): IoU pairs counted by an
        # un-timed instrumented launch x 15 FP ops per pair over the NMS-phase time the kernel stamps.
        # The pair test runs in float32 (rigorous screen) with the float64 arithmetic of the reference only for the
        # pairs inside the error bound, so the float32 peak is the denominator.
        roof_alu = None
        if phases and phases.get("nms_stripes"):
            pairs = counts["pairs_per_launch"]
            gops = pairs * 15 / (phases["nms_stripes"] * 1e-6) / 1e9
            roof_alu = {"phase": "NMS stripes of proposal_fused_kernel", "bound": "alu-f32", "achieved": gops,
                        "peak": alu["f32"], "unit": "Gop/s", "frac": gops / alu["f32"],
                        "pairs_evaluated_per_launch": pairs, "pairs_source": "counted by an instrumented launch (mxr_proposal_count_pairs)",
                        "float64_pairs_per_launch": counts["float64_pairs_per_launch"], "ops_per_pair": 15,
                        "full_triangle_pairs_per_launch": B * 17997000, "nms_phase_us": phases["nms_stripes"],
                        "peak_source": "mxr_bench_alu_peak (independent FADD/FMUL streams, -fmad=false)",
                        "f64_peak_gops": alu["f64"], "frac_of_f64_peak": gops / alu["f64"],
                        "note": "the stripes evaluate %.1f%% of the full triangle's pairs (windows of the score order are "
                                "materialised only until post_nms_top_n boxes are kept); a lower pair count lowers this "
                                "fraction and the step time together" % (100.0 * pairs / (B * 17997000.0))}
            if sat:
                # same pairs per image, whole step time of the saturating batch (all phases, not only the NMS stripes)
                sat_gops = pairs / float(B) * sat["batch"] * 15 / (sat["ms_per_step"] * 1e-3) / 1e9
                roof_alu["saturating_batch_frac_whole_step"] = sat_gops / alu["f32"]
        # issue-slot view: warp instructions per launch (from the committed ncu capture) over the step time against
        # 148 SMs x 4 schedulers x SM clock
        roof_issue = None
        if ncu and ncu.get("workload") == WORKLOAD and ncu.get("warp_instructions"):
            clk = (clocks or {}).get("sm_mhz") or 1965.0
            peak_issue = 148 * 4 * clk * 1e6
            roof_issue = {"warp_instructions_per_launch": ncu["warp_instructions"],
                          "achieved_per_s": ncu["warp_instructions"] / (ms_step * 1e-3), "peak_per_s": peak_issue,
                          "frac": ncu["warp_instructions"] / (ms_step * 1e-3) / peak_issue,
                          "source": "profiles/r02_ncu_metrics.json"}
            if sat and ncu.get("warp_instructions_saturating"):
                roof_issue["saturating_batch_frac"] = ncu["warp_instructions_saturating"] / (sat["ms_per_step"] * 1e-3) / peak_issue
        kind = reference_kind()
        n_cpu = max(host_cores(), min(12 * host_cores(), 96))
        cpu_block = cpu_baseline_block(n_cpu, kind)
        try:
            configs = secondary_configs(lib, _lib, dev, peak, alu)
        except Exception as ex:  # a secondary table must not take the contract line down
            configs = {"error": repr(ex)}
        line = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_per_gpu": B, "gpu_preconditioning_ms": PRECONDITION_MS,
                       "l2_policy": "inputs rotate through a %d-batch pool of %.0f MB (> 126 MB L2)" % (pool, pool * per_batch / 1e6),
                       "arithmetic": "box decode, IoU and outputs in the reference's float64 (exp in float32); float32 screens "
                                     "with rigorous error bounds in front of the min-size and IoU tests"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": 1e3 * e2e_s / args.steps,
                    "api": "mxr_proposal_submit_host / mxr_proposal_wait_host, two slots, inputs in the reference's layout "
                           "(cls_prob, bbox_pred, im_info as separate pinned arrays)",
                    "bound": "host link (H2D)", "h2d_gbs": h2d * args.steps / e2e_s / 1e9,
                    "host_link": {"plain_pinned_copy_64MiB_gbs": link_big, "plain_pinned_copy_step_bytes_gbs": link_step,
                                  "frac_of_link": (h2d * args.steps / e2e_s / 1e9) / link_big if link_big else None,
                                  "per_rank": [{"h2d_gbs": a_, "plain_copy_64MiB_gbs": b__, "plain_copy_step_bytes_gbs": c_}
                                               for a_, b__, c_ in per_rank],
                                  "note": "plain copies run on all ranks at once: what the host's memory and PCIe complex "
                                          "give N concurrent streams"},
                    "caller_packed": {"value": world * B * args.steps / e2e_packed_s, "ms_per_step": 1e3 * e2e_packed_s / args.steps,
                                      "h2d_gbs": h2d * args.steps / e2e_packed_s / 1e9,
                                      "api": "mxr_proposal_submit_host_packed (fg scores | deltas | im_info in one pinned block)"},
                    "one_call_at_a_time": {"value": e2e_sync_value, "ms_per_step": 1e3 * e2e_sync_s / args.steps,
                                           "api": "mxr_proposal_forward_host"}},
            "gpu_launches": launches_per_step * args.steps,
            "roofline": roof,
            "roofline_alu": roof_alu,
            "roofline_issue": roof_issue,
            "saturating_batch": sat,
            "single_image": lat1,
            "two_batches_in_flight": two_in_flight,
            "kernels": per_kernel,
            "fused_kernel_phases_us": phases,
            "configs": configs,
            "cpu_baseline": cpu_block,
            "parity_spot_check": ok, "allgather_checksum": checksum,
        }
        emit(line)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


_REAL_STDOUT = None


def claim_stdout():
    """stdout carries exactly ONE JSON line.  Libraries write there too (NCCL prints its version banner on stdout),
    so file descriptor 1 is pointed at stderr for the rest of the run and the line goes to the saved descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    run_gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
