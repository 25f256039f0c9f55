"""CPU tests of the rs.py drop-in's host logic: window generator, tile plans, frame sharding (gloo, world 2)."""
import os

import numpy as np
import pytest
import torch

from image_super_resolution_b200 import rs
from oracle import modelpt_oracle as mo


def test_sliding_window_matches_reference_semantics():
    img = torch.arange(3 * 20 * 29, dtype=torch.int32).reshape(3, 20, 29)
    ours = [(tuple(s), x, y, w.clone()) for s, x, y, w in rs.sliding_window(img, 8)]
    ref = [(tuple(s), x, y, w.clone()) for s, x, y, w in mo.sliding_window(img, 8)]
    assert len(ours) == len(ref) == 3 * 4 == rs.total_inter_sliding_window(img, 8)
    for a, b in zip(ours, ref):
        assert a[:3] == b[:3] and torch.equal(a[3], b[3])
    # window larger than the image: step is clamped, one (ragged) window
    one = list(rs.sliding_window(img, 96))
    assert len(one) == 1 and one[0][3].shape == img.shape


@pytest.mark.parametrize("H,W,th,tw", [(64, 64, 32, 32), (100, 37, 48, 16), (5, 7, 96, 96)])
def test_plan_tiles_covers_image_once_with_clipped_halo(H, W, th, tw):
    cover = np.zeros((H, W), dtype=np.int32)
    for y, x, h, w, y0, x0, y1, x1 in rs.plan_tiles(H, W, th, tw, 25):
        cover[y:y + h, x:x + w] += 1
        assert 0 <= y0 <= y and 0 <= x0 <= x and y + h <= y1 <= H and x + w <= x1 <= W
        assert y - y0 == min(25, y) and x - x0 == min(25, x)
    assert (cover == 1).all()


def test_shard_round_robin():
    got = sorted(i for r in range(3) for i in rs.shard(10, r, 3))
    assert got == list(range(10))
    assert list(rs.shard(10, 1, 3)) == [1, 4, 7]
    assert list(rs.shard(2, 3, 8)) == []  # more GPUs than units: idle rank


def test_byte_sink_signature():
    s = rs.ByteSink(keep=True)
    s.writeFrame(np.zeros((4, 6, 3), dtype=np.uint8))
    assert s.frames == 1 and s.bytes == 72 and s.kept[0].shape == (4, 6, 3)


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(5)
    frames = torch.randint(0, 256, (5, 3, 12, 10), dtype=torch.uint8, generator=g)  # same on every rank
    sd = mo.load_weights()
    mine = {i: mo.forward_u8(frames[i:i + 1], sd)[0] for i in rs.shard(len(frames), rank, world)}
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    out = torch.stack([{k: v for d in gathered for k, v in d.items()}[i] for i in range(len(frames))])
    if rank == 0:
        q.put(out.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_frame_sharding_world2_gloo_equals_single_process():
    """N>1 path on CPU: two gloo ranks each process their shard (oracle stands in for the GPU model) and the
    gathered result, in frame order, equals the single-process result: no collective on the data path."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 1000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    g = torch.Generator().manual_seed(5)
    frames = torch.randint(0, 256, (5, 3, 12, 10), dtype=torch.uint8, generator=g)
    want = mo.forward_u8(frames).numpy()
    assert np.array_equal(got, want)
