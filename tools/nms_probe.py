"""Times mxr_nms_batched on device-resident dets for the workloads of the NMS callers (per-class loop of
rcnn/tester.py:50-56, one long list, the proposal op's 6000-box list) with the fused NMS kernel (nms_fused.cuh) and
with the first-generation cluster kernel (MXRCNN_B200_V1=1), checks both against the oracle, prints one JSON line per
workload.  Usage: python tools/nms_probe.py [--cluster 2,4,8,16]"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cluster", default="")
    ap.add_argument("--reps", type=int, default=200)
    args = ap.parse_args()
    import torch
    from mxrcnn_b200 import _lib
    from oracle import restate, synth
    lib = _lib.load()
    handles = {}
    os.environ.pop("MXRCNN_B200_CLUSTER", None)
    os.environ["MXRCNN_B200_V1"] = "1"
    handles["v1"] = _lib.Handle(0)
    del os.environ["MXRCNN_B200_V1"]
    handles["fused"] = _lib.Handle(0)
    for c in [int(x) for x in args.cluster.split(",") if x]:
        os.environ["MXRCNN_B200_CLUSTER"] = str(c)
        handles["fused_c%d" % c] = _lib.Handle(0)
        del os.environ["MXRCNN_B200_CLUSTER"]
    work = [("per_class_20x300_f32", [300] * 20, np.float32, 0.3), ("per_class_20x300_f64", [300] * 20, np.float64, 0.3),
            ("one_2000_f32", [2000], np.float32, 0.7), ("one_2000_f64", [2000], np.float64, 0.7),
            ("one_6000_f64", [6000], np.float64, 0.7), ("eight_6000_f32", [6000] * 8, np.float32, 0.7),
            ("one_300_f32", [300], np.float32, 0.3), ("eighty_100_f32", [100] * 80, np.float32, 0.3)]
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream().cuda_stream
    for name, sizes, dt, th in work:
        dets = [synth.class_dets(900 + i, n=n, dtype=dt, clusters=max(2, n // 40)) for i, n in enumerate(sizes)]
        want = [[int(i) for i in restate.nms(d, th)] for d in dets[:4]]
        offsets = np.zeros((len(dets) + 1,), dtype=np.int32)
        offsets[1:] = np.cumsum(sizes)
        flat = torch.from_numpy(np.concatenate(dets, axis=0)).to(dev)
        keep = torch.empty((int(offsets[-1]),), dtype=torch.int32, device=dev)
        nk = torch.zeros((len(dets), 2), dtype=torch.int32, device=dev)
        row = {"workload": name, "kept_first": len(want[0])}
        for hn, h in handles.items():
            def call():
                _lib.check(lib.mxr_nms_batched(h.ptr, stream, flat.data_ptr(), offsets.ctypes.data_as(_lib.c_int_p),
                                               len(dets), 1 if dt == np.float64 else 0, float(th), 0, keep.data_ptr(),
                                               nk.data_ptr()))
            for _ in range(5):
                call()
            torch.cuda.synchronize()
            kh, nh = keep.cpu().numpy(), nk.cpu().numpy()
            ok = all([int(i) for i in kh[offsets[c]:offsets[c] + nh[c, 0]]] == want[c] for c in range(len(want)))
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(args.reps):
                call()
            b.record()
            torch.cuda.synchronize()
            row[hn + "_us"] = round(a.elapsed_time(b) * 1e3 / args.reps, 1)
            row[hn + "_ok"] = bool(ok)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
