"""How much of a proposal step at configs[1] is the cold read of its inputs?  Times the same launches over input pools
of 2 batches (L2-resident), 39 batches (161 MB, bench.py's pool) and 96 batches (397 MB), for 20 and 200 steps.
Usage: python tools/l2_probe.py"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import bench
    from mxrcnn_b200 import _lib
    from oracle import synth
    lib = _lib.load()
    dev = torch.device("cuda", 0)
    cfg = synth.CONFIGS["voc_test"]
    B, A = 8, 9
    base = 8
    cs, bs, infs = [], [], []
    for i in range(base):
        c, b_, inf = synth.proposal_batch([1000003 + i * B + j for j in range(B)], cfg)
        cs.append(torch.from_numpy(c))
        bs.append(torch.from_numpy(b_))
        infs.append(torch.from_numpy(inf))
    c8, b8, i8 = torch.stack(cs).to(dev), torch.stack(bs).to(dev), torch.stack(infs).to(dev)
    h = _lib.Handle(0)
    spin = torch.randn((4096, 4096), device=dev)
    for pool in (2, 39, 96):
        reps = (pool + base - 1) // base
        cls_d = c8.repeat(reps, 1, 1, 1, 1)[:pool].contiguous()
        box_d = b8.repeat(reps, 1, 1, 1, 1)[:pool].contiguous()
        info_d = i8.repeat(reps, 1, 1)[:pool].contiguous()
        run = bench.ProposalRunner(lib, _lib, h, dev, cfg, B, cls_d, box_d, info_d)
        row = {"pool_batches": pool, "pool_mb": round(pool * B * 6 * A * cfg["H"] * cfg["W"] * 4 / 1e6)}
        for steps in (20, 200):
            ts = []
            for trial in range(7):
                for _ in range(6):
                    spin @ spin  # clocks up, L2 filled with something else
                torch.cuda.synchronize()
                with torch.cuda.stream(run.stream):
                    run.timed(trial * 7, 5)
                    ts.append(run.timed(trial * 7 + 5, steps) * 1e3)
            row["us_per_step_%d" % steps] = round(float(np.median(ts)), 2)
            row["min_%d" % steps] = round(float(np.min(ts)), 2)
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
