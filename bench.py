#!/usr/bin/env python
"""Benchmark of the hot path: RPN Proposal+NMS images/s (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one proposal-op pass over a batch of 8 synthetic 600x1000 images per GPU
(BASELINE.json configs[1]: 38x63 map, 9 anchors, pre-NMS 6000, NMS 0.7, post-NMS 300).
N > 1 is launched by torchrun, one rank per GPU; images shard across ranks with no
collective in the timed path ("scaling": "weak").  Rank 0 prints ONE JSON line.

  value        images/s, inputs resident in HBM, CUDA events on the launching stream,
               barrier + synchronize on both sides, max over ranks.
  e2e          images/s through the C-ABI host entry point (mxr_proposal_forward_host):
               pinned host inputs, H2D + kernels + D2H inside the timed region.
  roofline     dominant kernel (by measured share of the step) against MEASURED_PEAKS.json.
  cpu_baseline oracle/restate.py (numpy restatement of the reference) on the host cores.
  --impl reference  times only that CPU restatement (the reference is Python and is not
               mounted on the GPU box; oracle/restate.py follows it line for line and is pinned
               by tests/golden fixtures generated from the real reference).
"""
import argparse
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


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# -----------------------------------------------------------------------------------------
# CPU arm: oracle/restate.py on the host cores
# -----------------------------------------------------------------------------------------
def _cpu_worker(args):
    seeds, cfg_name = args
    from oracle import restate, synth
    cfg = synth.CONFIGS[cfg_name]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    data = [synth.proposal_inputs(s, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"]) for s in seeds]
    t0 = time.perf_counter()
    for cls, box, info in data:
        restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"],
                                cfg["post"], cfg["thresh"], cfg["min_size"])
    return time.perf_counter() - t0


def cpu_throughput(n_images, procs, seed0=900000):
    """images/s of the numpy restatement with `procs` worker processes, each looping over its share
    of the images (per-image independence is the reference's only parallelism)."""
    import multiprocessing as mp
    for v in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
        os.environ[v] = "1"
    seeds = [seed0 + i for i in range(n_images)]
    shares = [seeds[i::procs] for i in range(procs)]
    shares = [s for s in shares if s]
    if len(shares) == 1:
        t0 = time.perf_counter()
        _cpu_worker((shares[0], CONFIG))
        return n_images / (time.perf_counter() - t0)
    ctx = mp.get_context("fork")
    with ctx.Pool(len(shares)) as pool:
        pool.map(_cpu_worker, [([shares[0][0]], CONFIG)] * len(shares))  # warm the workers (imports)
        t0 = time.perf_counter()
        pool.map(_cpu_worker, [(s, CONFIG) for s in shares])
        wall = time.perf_counter() - t0
    return n_images / wall


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    procs = host_cores()
    per_step = BATCH  # one step = one batch of 8 images, like the GPU arm at N=1
    # bounded: each step is BATCH images spread over the cores (~0.1 s of CPU work per image)
    steps = max(1, min(args.steps, 40))
    n = 0
    import multiprocessing as mp
    for v in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
        os.environ[v] = "1"
    ctx = mp.get_context("fork")
    shares_n = min(procs, per_step)
    with ctx.Pool(shares_n) as pool:
        for _ in range(max(1, min(args.warmup, 2))):  # warm-up: imports, page-in
            pool.map(_cpu_worker, [([1], CONFIG)] * shares_n)
        t0 = time.perf_counter()
        for st in range(steps):
            seeds = [800000 + st * per_step + i for i in range(per_step)]
            pool.map(_cpu_worker, [(seeds[i::shares_n], CONFIG) for i in range(shares_n)])
            n += per_step
        wall = time.perf_counter() - t0
    v = n / wall
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": min(args.warmup, 2), "ms_per_step": 1e3 * wall / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "voc_test 38x63x9 anchors, pre 6000, nms 0.7, post 300, batch 8 per step", "batch": BATCH},
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": shares_n, "kind": "port",
                         "sample": "%d steps x %d images, oracle/restate.py (numpy restatement of rcnn/rpn/proposal.py "
                                   "+ helper/processing/nms.py), one worker process per image" % (steps, per_step)},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
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
        self.index, self.rows, self.proc = index, [], None

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
        for r in self.rows:
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
def algorithmic_bytes(cfg, B):
    """SURVEY.md 8(d) per-image figures x images per launch (DESIGN.md section 'Kernels')."""
    A = len(cfg["scales"]) * len(cfg["ratios"])
    N = A * cfg["H"] * cfg["W"]
    k = min(cfg["pre"], N) if cfg["pre"] > 0 else N
    post = cfg["post"]
    return {
        # whole op, mandatory I/O: fg scores 4N + deltas 16N + im_info 12 + rois 20*post + scores 4*post
        "op": B * (20 * N + 12 + 24 * post),
        # decode/filter: reads fg scores 4N + deltas 16N + im_info 12, writes one 4-byte key per anchor
        "decode_filter_kernel": B * (24 * N + 12),
        # select+sort+gather: reads keys 4N, gathers 5 floats per selected box, writes box/area/score/index (48 B)
        "topk_sort_kernel": B * (4 * N + 68 * k),
        # fused kernel = the whole op: its algorithmic bytes are the op's mandatory I/O
        "proposal_cluster_kernel": B * (20 * N + 12 + 24 * post),
    }


def run_gpu_arm(args, rank, world, local_rank):
    import ctypes
    import torch
    from oracle import synth
    from mxrcnn_b200 import _lib
    from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        # stdout carries the ONE JSON line: whatever NCCL logs (its version banner at any NCCL_DEBUG level, INFO
        # traces when the caller asks for them) goes to a per-process file instead
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/mxrcnn_b200_nccl_%h_%p.log")
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=dev)
    cfg = synth.CONFIGS[CONFIG]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    B = BATCH
    post = cfg["post"]
    lib = _lib.load()
    h = _lib.Handle(local_rank)
    anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
    anc = anchors.ctypes.data_as(_lib.c_double_p)
    p = _lib.ProposalParams(B, A, cfg["H"], cfg["W"], 0, 0, float(cfg["feat_stride"]), cfg["pre"], post,
                            cfg["thresh"], float(cfg["min_size"]), 0)
    # ---- synthetic input pool: larger than L2 (126 MB), rotated so no step re-reads cached inputs ----
    per_batch = B * (2 * A + 4 * A) * cfg["H"] * cfg["W"] * 4
    pool = max(4, int(np.ceil(160e6 / per_batch)))
    log("[bench] rank %d: building %d input batches (%.0f MB > L2)" % (rank, pool, pool * per_batch / 1e6))
    cls_h = torch.empty((pool, B, 2 * A, cfg["H"], cfg["W"]), dtype=torch.float32).pin_memory()
    box_h = torch.empty((pool, B, 4 * A, cfg["H"], cfg["W"]), dtype=torch.float32).pin_memory()
    info_h = torch.empty((pool, B, 3), dtype=torch.float32).pin_memory()
    for i in range(pool):
        seeds = [1000003 * (rank + 1) + i * B + j for j in range(B)]
        c, b_, inf = synth.proposal_batch(seeds, cfg)
        cls_h[i].copy_(torch.from_numpy(c))
        box_h[i].copy_(torch.from_numpy(b_))
        info_h[i].copy_(torch.from_numpy(inf))
    cls_d, box_d, info_d = cls_h.to(dev), box_h.to(dev), info_h.to(dev)
    rois = torch.zeros((B * post, 5), dtype=torch.float32, device=dev)
    scores = torch.zeros((B * post,), dtype=torch.float32, device=dev)
    keep = torch.zeros((B, post), dtype=torch.int32, device=dev)
    nk = torch.zeros((B, 2), dtype=torch.int32, device=dev)
    stream = torch.cuda.Stream(dev)  # a real (non-default) stream: kernels and timing events share it
    sp = int(stream.cuda_stream)
    torch.cuda.synchronize(dev)

    def step(i):
        j = i % pool
        st = lib.mxr_proposal_forward(h.ptr, sp, ctypes.byref(p), cls_d[j].data_ptr(), box_d[j].data_ptr(),
                                      info_d[j].data_ptr(), anc, post, rois.data_ptr(), scores.data_ptr(),
                                      keep.data_ptr(), None, nk.data_ptr())
        if st != 0:
            _lib.check(st)

    def barrier():
        torch.cuda.synchronize(dev)
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    lib.mxr_reserve_proposal(h.ptr, ctypes.byref(p))
    for i in range(max(args.warmup, 3)):
        step(i)
    torch.cuda.synchronize(dev)
    assert int(nk[:, 0].min()) == post and int(nk[:, 1].max()) == 0, nk
    launches_per_step = h.last_launch_count()

    # ---- timed region: device-resident inputs ----
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record(stream)
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    value = world * B * args.steps / (ms * 1e-3)

    # ---- same K steps with two batches in flight (not the headline): one step occupies 8 clusters x 8 SMs of 148,
    #      so a second handle (own workspace, own stream, own outputs) runs the odd steps next to the even ones ----
    h_b = _lib.Handle(local_rank)
    lib.mxr_reserve_proposal(h_b.ptr, ctypes.byref(p))
    stream_b = torch.cuda.Stream(dev)
    outs_b = (torch.zeros_like(rois), torch.zeros_like(scores), torch.zeros_like(keep), torch.zeros_like(nk))

    def step_pair(i):
        j = i % pool
        if i % 2 == 0:
            st = lib.mxr_proposal_forward(h.ptr, sp, ctypes.byref(p), cls_d[j].data_ptr(), box_d[j].data_ptr(),
                                          info_d[j].data_ptr(), anc, post, rois.data_ptr(), scores.data_ptr(),
                                          keep.data_ptr(), None, nk.data_ptr())
        else:
            st = lib.mxr_proposal_forward(h_b.ptr, int(stream_b.cuda_stream), ctypes.byref(p), cls_d[j].data_ptr(),
                                          box_d[j].data_ptr(), info_d[j].data_ptr(), anc, post, outs_b[0].data_ptr(),
                                          outs_b[1].data_ptr(), outs_b[2].data_ptr(), None, outs_b[3].data_ptr())
        if st != 0:
            _lib.check(st)

    for i in range(4):
        step_pair(i)
    barrier()
    f0, f1, fb = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    f0.record(stream)
    stream_b.wait_event(f0)
    for i in range(args.steps):
        step_pair(args.warmup + i)
    fb.record(stream_b)
    stream.wait_event(fb)
    f1.record(stream)
    barrier()
    ms_pair = max_over_ranks(f0.elapsed_time(f1))
    assert int(outs_b[3][:, 0].min()) == post
    two_in_flight = {"value": world * B * args.steps / (ms_pair * 1e-3), "unit": "images/s",
                     "ms_per_step": ms_pair / args.steps,
                     "how": "steps alternate between two handles / streams; same kernels, same K steps"}
    h_b.close()

    # ---- e2e: host buffers through the C-ABI host entry point ----
    cls_np, box_np, info_np = cls_h.numpy(), box_h.numpy(), info_h.numpy()
    o_rois = np.empty((B * post, 5), np.float32)
    o_scores = np.empty((B * post,), np.float32)
    o_keep = np.empty((B, post), np.int32)
    o_nk = np.zeros((B, 2), np.int32)

    def step_host(i):
        j = i % pool
        st = lib.mxr_proposal_forward_host(h.ptr, ctypes.byref(p), _lib.hptr(cls_np[j]), _lib.hptr(box_np[j]),
                                           _lib.hptr(info_np[j]), anc, post, _lib.hptr(o_rois), _lib.hptr(o_scores),
                                           _lib.hptr(o_keep), None, _lib.hptr(o_nk))
        if st != 0:
            _lib.check(st)

    # pipelined form of the same entry point: two slots, so the H2D copy of batch i+1 overlaps the kernels
    # of batch i; every step still copies its own inputs in and its own results out inside the timed region
    outs = [(np.empty((B * post, 5), np.float32), np.empty((B * post,), np.float32), np.empty((B, post), np.int32),
             np.zeros((B, 2), np.int32)) for _ in range(2)]

    def submit(i):
        j = i % pool
        st = lib.mxr_proposal_submit_host(h.ptr, i % 2, ctypes.byref(p), _lib.hptr(cls_np[j]), _lib.hptr(box_np[j]),
                                          _lib.hptr(info_np[j]), anc, post)
        if st != 0:
            _lib.check(st)

    def wait(i):
        o = outs[i % 2]
        st = lib.mxr_proposal_wait_host(h.ptr, i % 2, _lib.hptr(o[0]), _lib.hptr(o[1]), _lib.hptr(o[2]), None,
                                        _lib.hptr(o[3]))
        if st != 0:
            _lib.check(st)
        assert o[3][0, 0] == post

    def run_pipelined(first, count):
        submit(first)
        for i in range(first + 1, first + count):
            submit(i)
            wait(i - 1)
        wait(first + count - 1)

    for i in range(max(args.warmup, 3)):
        step_host(i)
    run_pipelined(0, max(args.warmup, 4))
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step_host(args.warmup + i)
    torch.cuda.synchronize(dev)
    e2e_sync_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    t0 = time.perf_counter()
    run_pipelined(args.warmup, args.steps)
    torch.cuda.synchronize(dev)
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    # both forms processed the same last batch: identical keep lists
    assert np.array_equal(outs[(args.warmup + args.steps - 1) % 2][2], o_keep)
    e2e_sync_value = world * B * args.steps / e2e_sync_s
    clocks = sampler.stop() if rank == 0 else None
    e2e_value = world * B * args.steps / e2e_s
    N = A * cfg["H"] * cfg["W"]
    h2d = B * (4 * N + 16 * N + 12)  # fg half of cls_prob + deltas + im_info (the bg half is never copied)
    d2h = B * post * (20 + 4 + 4 + 4) + B * 8
    # what the host link gives a plain pinned -> device copy of one step's input bytes (the e2e bound on this box)
    link_gbs = None
    if rank == 0:
        pin = torch.empty(h2d, dtype=torch.uint8).pin_memory()
        dst = torch.empty(h2d, dtype=torch.uint8, device=dev)
        with torch.cuda.stream(stream):
            for _ in range(5):
                dst.copy_(pin, non_blocking=True)
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record(stream)
            for _ in range(50):
                dst.copy_(pin, non_blocking=True)
            c1.record(stream)
        stream.synchronize()
        link_gbs = h2d * 50 / (c0.elapsed_time(c1) * 1e-3) / 1e9
        del pin, dst

    # ---- per-kernel profile (separate pass, CUDA events around every launch on the launching stream) ----
    prof_steps = min(args.steps, 50)
    h.profile_enable(True)
    for i in range(prof_steps):
        step(i)
    prof = h.profile_read()
    h.profile_enable(False)
    phases = h.proposal_phase_times()
    phase_mean = {k_: float(np.mean([p_[k_] for p_ in phases])) for k_ in phases[0]} if phases else None

    # ---- measured ALU peaks (in-repo micro-benchmark, no FMA contraction) for the NMS roofline ----
    alu = {}
    for nm, code in (("f64", 1), ("f32", 0)):
        gv = ctypes.c_double(0)
        _lib.check(lib.mxr_bench_alu_peak(h.ptr, code, ctypes.byref(gv)))
        alu[nm] = gv.value

    # ---- saturating batch (not the headline): 256 images in one launch = 132 MB of inputs (> L2),
    #      which the library runs as 4 waves of 2-CTA clusters; shows the throughput asymptote of the same kernel ----
    sat = None
    if rank == 0:
        SB = 256
        reps = (SB + pool * B - 1) // (pool * B)
        cls_s = cls_d.reshape(pool * B, 2 * A, cfg["H"], cfg["W"]).repeat(reps, 1, 1, 1)[:SB].contiguous()
        box_s = box_d.reshape(pool * B, 4 * A, cfg["H"], cfg["W"]).repeat(reps, 1, 1, 1)[:SB].contiguous()
        info_s = info_d.reshape(pool * B, 3).repeat(reps, 1)[:SB].contiguous()
        ps = _lib.ProposalParams(SB, A, cfg["H"], cfg["W"], 0, 0, float(cfg["feat_stride"]), cfg["pre"], post,
                                 cfg["thresh"], float(cfg["min_size"]), 0)
        rois_s = torch.zeros((SB * post, 5), dtype=torch.float32, device=dev)
        nk_s = torch.zeros((SB, 2), dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)

        def step_sat():
            st = lib.mxr_proposal_forward(h.ptr, sp, ctypes.byref(ps), cls_s.data_ptr(), box_s.data_ptr(),
                                          info_s.data_ptr(), anc, post, rois_s.data_ptr(), None, None, None,
                                          nk_s.data_ptr())
            if st != 0:
                _lib.check(st)

        for _ in range(3):
            step_sat()
        torch.cuda.synchronize(dev)
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record(stream)
        nsat = 20
        for _ in range(nsat):
            step_sat()
        s1.record(stream)
        torch.cuda.synchronize(dev)
        sat_ms = s0.elapsed_time(s1) / nsat
        assert int(nk_s[:, 0].min()) == post
        sat = {"batch": SB, "value": SB / (sat_ms * 1e-3), "unit": "images/s", "ms_per_step": sat_ms,
               "op_hbm_gbs": SB * (20 * A * cfg["H"] * cfg["W"] + 12 + 24 * post) / (sat_ms * 1e-3) / 1e9}
        del cls_s, box_s, info_s, rois_s

    # ---- parity spot check of the last batch against the oracle (not timed) ----
    ok = None
    if rank == 0:
        from oracle import restate
        step(0)
        torch.cuda.synchronize(dev)
        c0, b0, i0 = cls_np[0][:1], box_np[0][:1], info_np[0][:1]
        st_ = restate.proposal_stages(c0, b0, i0, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"],
                                      post, cfg["thresh"], cfg["min_size"], nms_stop_after=post)
        ok = bool(np.array_equal(keep[0].cpu().numpy(), st_["keep"]))

    # ---- final all-gather of the per-image outputs (harness only; not in the timed path) ----
    checksum = None
    if dist is not None:
        from mxrcnn_b200 import shard
        # rank r owns global images r, r+world, ...: gather the keep lists of the last global batch
        full = shard.all_gather_outputs(keep, world * B)
        checksum = shard.checksum(full.cpu().numpy())

    if rank == 0:
        peak, peak_src = measured_peaks()
        alg = algorithmic_bytes(cfg, B)
        per_kernel = {}
        step_ms = 0.0
        for name, (tot, cnt) in prof.items():
            per_kernel[name] = {"ms_per_step": tot / prof_steps, "launches_per_step": cnt / prof_steps,
                                "us_per_launch": 1e3 * tot / cnt}
            step_ms += tot / prof_steps
        for name in per_kernel:
            per_kernel[name]["share"] = per_kernel[name]["ms_per_step"] / step_ms if step_ms else None
        dom = max(per_kernel, key=lambda k_: per_kernel[k_]["ms_per_step"]) if per_kernel else None
        roof = None
        if dom:
            d = per_kernel[dom]
            per_launch_s = d["us_per_launch"] * 1e-6
            bytes_launch = alg.get(dom, alg["op"]) / max(1.0, 1.0)
            ach = bytes_launch / per_launch_s / 1e9
            # dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` capture of this kernel at this
            # configuration (profiles/r01h_cluster_kernel_ncu.md); not re-measured here (never time under a profiler)
            traffic = 3893248 if (dom == "proposal_cluster_kernel" and B == 8 and CONFIG == "voc_test") else None
            roof = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_launch,
                    "us_per_launch": d["us_per_launch"], "share_of_step": d["share"]}
        # ALU view of the fused kernel's NMS phase (BASELINE.json: "NMS by SM FP32/ALU-pipe utilisation"):
        # IoU pairs actually evaluated (256-box stripes: in-stripe triangle + kept-so-far x stripe, the latter
        # bounded by post x 256) x 15 FP ops per pair over the measured NMS-phase time
        roof_alu = None
        if phase_mean and phase_mean.get("nms_stripes"):
            stripes = phase_mean["stripes_walked"]
            pairs = B * (stripes * (256 * 255 // 2) + max(0.0, stripes - 1) * 0.5 * post * 256)
            gflops = pairs * 15 / (phase_mean["nms_stripes"] * 1e-6) / 1e9
            roof_alu = {"phase": "nms_stripes of proposal_cluster_kernel", "bound": "alu-f64", "achieved": gflops,
                        "peak": alu["f64"], "unit": "Gop/s", "frac": gflops / alu["f64"],
                        "pairs_evaluated_per_launch": pairs, "ops_per_pair": 15,
                        "full_triangle_pairs_per_launch": B * 17997000,
                        "peak_source": "mxr_bench_alu_peak (independent DADD/DMUL streams, -fmad=false)",
                        "f32_peak_gops": alu["f32"]}
        # whole-op HBM fraction (SURVEY.md 8d mandatory I/O per image)
        op_gbs = alg["op"] / (ms * 1e-3 / args.steps) / 1e9
        cores = host_cores()
        n_cpu = max(cores, min(16 * cores, 96))
        cpu_v = cpu_throughput(n_cpu, cores)
        line = {
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "voc_test: 38x63 map x 9 anchors (21546), pre-NMS 6000, NMS 0.7, post-NMS 300, "
                                   "batch 8 images per GPU per step (BASELINE.json configs[1])",
                       "batch_per_gpu": B, "l2_policy": "inputs rotate through a %d-batch pool of %.0f MB (> 126 MB L2)"
                                                        % (pool, pool * per_batch / 1e6)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": 1e3 * e2e_s / args.steps,
                    "api": "mxr_proposal_submit_host / mxr_proposal_wait_host, two slots (pinned host buffers)",
                    "bound": "host link (H2D)", "h2d_gbs": h2d / (e2e_s / args.steps) / 1e9,
                    "h2d_link_gbs_plain_copy": link_gbs,
                    "one_call_at_a_time": {"value": e2e_sync_value, "ms_per_step": 1e3 * e2e_sync_s / args.steps,
                                           "api": "mxr_proposal_forward_host"}},
            "gpu_launches": launches_per_step * args.steps,
            "roofline": roof,
            "roofline_alu": roof_alu,
            "saturating_batch": sat,
            "two_batches_in_flight": two_in_flight,
            "op_hbm": {"achieved": op_gbs, "peak": peak, "unit": "GB/s", "frac": op_gbs / peak,
                       "algorithmic_bytes_per_step": alg["op"]},
            "kernels": per_kernel,
            "fused_kernel_phases_us": phase_mean,
            "cpu_baseline": {"value": cpu_v, "unit": "images/s", "cores": cores, "kind": "port",
                             "sample": "%d images of the same workload, oracle/restate.py (numpy restatement), one "
                                       "process per core" % n_cpu},
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
