"""Secondary workloads of BASELINE.json (configs[2..4] + the reference's TRAIN / alternate-training
proposal settings): GPU time next to the numpy restatement on one host core.  bench.py keeps the one
contract line for configs[1]; this tool writes the table that DESIGN.md / profiles/ quote."""
import ctypes
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import restate, synth  # noqa: E402


def wall(fn, n, warm=2):
    for _ in range(warm):
        fn()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    return (time.perf_counter() - t0) / n


def proposal_row(name, B=8, steps=50):
    import torch
    from mxrcnn_b200 import _lib
    from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
    cfg = synth.CONFIGS[name]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    lib, h = _lib.load(), _lib.Handle(0)
    dev = torch.device("cuda:0")
    cls, box, info = synth.proposal_batch(list(range(500, 500 + B)), cfg)
    cd, bd, idv = (torch.from_numpy(x).to(dev) for x in (cls, box, info))
    post = cfg["post"]
    anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
    anc = anchors.ctypes.data_as(_lib.c_double_p)
    p = _lib.ProposalParams(B, A, cfg["H"], cfg["W"], 0, 0, float(cfg["feat_stride"]), cfg["pre"], post, cfg["thresh"],
                            float(cfg["min_size"]), 0)
    rois = torch.zeros((B * post, 5), dtype=torch.float32, device=dev)
    nk = torch.zeros((B, 2), dtype=torch.int32, device=dev)
    st = torch.cuda.Stream(dev)
    sp = int(st.cuda_stream)
    torch.cuda.synchronize()

    def step():
        _lib.check(lib.mxr_proposal_forward(h.ptr, sp, ctypes.byref(p), cd.data_ptr(), bd.data_ptr(), idv.data_ptr(), anc,
                                            post, rois.data_ptr(), None, None, None, nk.data_ptr()))

    for _ in range(5):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(steps):
        step()
    e1.record(st)
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / steps * 1e3
    h.profile_enable(True)
    for _ in range(3):
        step()
    h.profile_read()
    h.profile_enable(False)
    ph = h.proposal_phase_times()
    t_cpu = wall(lambda: restate.proposal_stages(cls[:1], box[:1], info[:1], cfg["feat_stride"], cfg["scales"], cfg["ratios"],
                                                 False, cfg["pre"], post, cfg["thresh"], cfg["min_size"],
                                                 rng=np.random.RandomState(0)), n=2, warm=1)
    return {"workload": "proposal %s (N=%d, pre %d, post %d), batch %d" % (name, A * cfg["H"] * cfg["W"], cfg["pre"], post, B),
            "gpu_us_per_call": us, "gpu_images_per_s": B / (us * 1e-6), "cpu_1core_ms_per_image": t_cpu * 1e3,
            "nkeep_min": int(nk[:, 0].min()), "stripes_walked": float(np.mean([x["stripes_walked"] for x in ph])) if ph else None}


def main():
    from mxrcnn_b200.helper.processing.nms import nms, nms_batched
    from mxrcnn_b200.rcnn.minibatch import assign_anchor
    from mxrcnn_b200.rcnn.rpn.proposal_target import _sample_rois
    rows = [proposal_row(n) for n in ("voc_test", "voc_train", "voc_alt", "wider")]
    info = np.array([[600, 1000, 1.6]], np.float32)
    gt64 = synth.gt_boxes(4, G=64)
    rows.append({"workload": "assign_anchor 38x63x9, 64 gt (configs[2]), host arrays in/out",
                 "gpu_us_per_call": wall(lambda: assign_anchor((1, 18, 38, 63), gt64, info), 20) * 1e6,
                 "cpu_1core_ms_per_image": wall(lambda: restate.assign_anchor((1, 18, 38, 63), gt64, info), 3, 1) * 1e3})
    gt8 = synth.gt_boxes(5, G=8)
    allr = np.vstack((synth.jittered_rois(5, gt8, R=2000), np.hstack((np.zeros((8, 1), np.float32), gt8[:, :4]))))
    rows.append({"workload": "_sample_rois 2008 rois x 8 gt, 128 sampled, 21 classes (configs[3]), host arrays",
                 "gpu_us_per_call": wall(lambda: _sample_rois(allr, gt8, 32, 128, 21, "TRAIN"), 20) * 1e6,
                 "cpu_1core_ms_per_image": wall(lambda: restate.sample_rois_e2e(allr, gt8, 32, 128, 21, "TRAIN"), 3, 1) * 1e3})
    per_class = [synth.class_dets(900 + c, n=300) for c in range(20)]
    rows.append({"workload": "per-class NMS 0.3, 20 classes x 300 dets f64 (tester.py loop), one batched launch",
                 "gpu_us_per_call": wall(lambda: nms_batched(per_class, 0.3), 20) * 1e6,
                 "cpu_1core_ms_per_image": wall(lambda: [restate.nms(d, 0.3) for d in per_class], 3, 1) * 1e3})
    d2000 = synth.class_dets(77, n=2000)
    rows.append({"workload": "nms 0.3, 2000 dets f64 (configs[4] test-time NMS), host arrays",
                 "gpu_us_per_call": wall(lambda: nms(d2000, 0.3), 20) * 1e6,
                 "cpu_1core_ms_per_image": wall(lambda: restate.nms(d2000, 0.3), 3, 1) * 1e3})
    from mxrcnn_b200.rcnn.tester import detections
    r, d, sc, shp, scl = synth.box_head_inputs(7)
    rows.append({"workload": "test-time box head 300 rois x 21 classes (detector.py:73-74 + tester.py:50-65), host arrays",
                 "gpu_us_per_call": wall(lambda: detections(sc, r, d, shp, scl), 20) * 1e6,
                 "cpu_1core_ms_per_image": wall(lambda: restate.box_head(r, d, sc, shp, scl), 3, 1) * 1e3})
    print(json.dumps(rows, indent=1))
    if len(sys.argv) > 1:  # markdown table for profiles/
        with open(sys.argv[1], "w") as f:
            f.write("| workload | GPU us / call | GPU images/s | numpy restatement ms / image (1 core) | note |\n|---|---|---|---|---|\n")
            for x in rows:
                note = ""
                if x.get("stripes_walked") is not None:
                    note = "%.0f stripes of 256 walked; nkeep >= %d" % (x["stripes_walked"], x["nkeep_min"])
                f.write("| %s | %.1f | %s | %.1f | %s |\n" % (
                    x["workload"], x["gpu_us_per_call"], ("%.0f" % x["gpu_images_per_s"]) if "gpu_images_per_s" in x else "-",
                    x["cpu_1core_ms_per_image"], note))


if __name__ == "__main__":
    main()
