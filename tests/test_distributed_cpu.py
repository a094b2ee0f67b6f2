"""CPU, world_size 2 over gloo: the image-sharding / all-gather logic of the multi-GPU path.

The per-image work in these processes is done by the ORACLE (no GPU here); what is tested is that
sharded results, gathered, equal the single-process results in global image order."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_images, out_dir):
    import torch
    import torch.distributed as dist
    from mxrcnn_b200 import shard
    from oracle import restate, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cfg = synth.CONFIGS["tiny"]
    mine = shard.shard_indices(n_images, rank, world)
    rows = []
    for i in mine:
        cls, box, info = synth.proposal_inputs(100 + i, H=cfg["H"], W=cfg["W"], A=9, im_info=cfg["im_info"])
        st = restate.proposal_stages(cls, box, info, 16, cfg["scales"], cfg["ratios"], False, cfg["pre"], cfg["post"],
                                     cfg["thresh"], cfg["min_size"], rng=np.random.RandomState(100 + i))
        rows.append(np.concatenate([st["rois"].ravel(), st["keep"].astype(np.float32)]))
    local = torch.from_numpy(np.stack(rows)) if rows else torch.zeros((0, cfg["post"] * 6))
    full = shard.all_gather_outputs(local, n_images)
    np.save(os.path.join(out_dir, "rank%d.npy" % rank), full.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_images", [6, 5])
def test_sharded_equals_single_process(tmp_path, n_images):
    import torch.multiprocessing as mp
    from oracle import restate, synth
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n_images, str(tmp_path)), nprocs=world, join=True)
    cfg = synth.CONFIGS["tiny"]
    want = []
    for i in range(n_images):
        cls, box, info = synth.proposal_inputs(100 + i, H=cfg["H"], W=cfg["W"], A=9, im_info=cfg["im_info"])
        st = restate.proposal_stages(cls, box, info, 16, cfg["scales"], cfg["ratios"], False, cfg["pre"], cfg["post"],
                                     cfg["thresh"], cfg["min_size"], rng=np.random.RandomState(100 + i))
        want.append(np.concatenate([st["rois"].ravel(), st["keep"].astype(np.float32)]))
    want = np.stack(want)
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), "rank%d.npy" % r))
        assert np.array_equal(got, want), "rank %d" % r


def test_shard_helpers():
    from mxrcnn_b200 import shard
    for n, w in [(8, 1), (8, 2), (8, 4), (8, 8), (7, 4), (3, 8)]:
        seen = sorted(i for r in range(w) for i in shard.shard_indices(n, r, w))
        assert seen == list(range(n))
        per = [[("img", i) for i in shard.shard_indices(n, r, w)] for r in range(w)]
        assert shard.unshard(per, n, w) == [("img", i) for i in range(n)]
    a = shard.checksum([[1, 2, 3], [4, 5]])
    assert a == shard.checksum([[1, 2, 3], [4, 5]]) and a != shard.checksum([[4, 5], [1, 2, 3]])
