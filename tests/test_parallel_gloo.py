"""Image sharding and the final detection gather (N > 1 path) on CPU: world size 2, gloo."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_images, q):
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "rp_parallel", os.path.join(ROOT, "instancesegmentation-jittor_b200", "parallel.py"))
    par = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(par)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, e = par.shard_range(n_images, rank, world)
    g = torch.Generator().manual_seed(17)
    per_img = torch.randint(0, 5, (n_images,), generator=g).tolist()  # detections per image (0 allowed)
    ids, cls, sc, bx, mk = [], [], [], [], []
    for i in range(b, e):
        gi = torch.Generator().manual_seed(100 + i)
        m = per_img[i]
        ids.append(torch.full((m,), i, dtype=torch.int64))
        cls.append(torch.randint(0, 80, (m,), generator=gi))
        sc.append(torch.rand(m, generator=gi))
        bx.append(torch.rand(m, 4, generator=gi))
        mk.append((torch.rand(m, 9, 13, generator=gi) > 0.5).to(torch.uint8))
    out = par.gather_detections(torch.cat(ids), torch.cat(cls), torch.cat(sc), torch.cat(bx),
                                torch.cat(mk))
    if rank == 0:
        q.put({k: v.numpy() for k, v in out.items()})
    else:
        assert out is None
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_covers_all_images():
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "rp_parallel", os.path.join(ROOT, "instancesegmentation-jittor_b200", "parallel.py"))
    par = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(par)
    for n, w in [(16, 8), (64, 8), (128, 8), (5, 2), (3, 4), (0, 2)]:
        seen = []
        for r in range(w):
            b, e = par.shard_range(n, r, w)
            seen += list(range(b, e))
        assert seen == list(range(n))
    m = (torch.rand(7, 5, 11) > 0.4).to(torch.uint8)
    assert torch.equal(par.unpack_masks(par.pack_masks(m), 5, 11), m)
    r = torch.tensor([[3., 1, 2, 3, 4], [4., 0, 0, 1, 1]])
    assert par.rebase_rois(r, 3)[:, 0].tolist() == [0.0, 1.0]


def test_gather_detections_world2_gloo():
    n_images, world = 7, 2
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_images, q)) for r in range(world)]
    [p.start() for p in procs]
    got = q.get()
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    g = torch.Generator().manual_seed(17)
    per_img = torch.randint(0, 5, (n_images,), generator=g).tolist()
    exp_ids = np.concatenate([np.full((per_img[i],), i, np.int64) for i in range(n_images)])
    np.testing.assert_array_equal(got["image_id"], exp_ids)
    sc, mk = [], []
    for i in range(n_images):
        gi = torch.Generator().manual_seed(100 + i)
        m = per_img[i]
        torch.randint(0, 80, (m,), generator=gi)
        sc.append(torch.rand(m, generator=gi))
        torch.rand(m, 4, generator=gi)
        mk.append((torch.rand(m, 9, 13, generator=gi) > 0.5).to(torch.uint8))
    np.testing.assert_array_equal(got["score"], torch.cat(sc).numpy())
    np.testing.assert_array_equal(got["mask"], torch.cat(mk).numpy())
