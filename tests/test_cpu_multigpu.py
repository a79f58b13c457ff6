"""CPU suite: the N > 1 host logic (sample sharding + one reduce of the sums per step) on
two gloo ranks."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pim_b200 import multigpu


def _fake_sample_image(sample, n=64):
    g = np.random.default_rng(sample)
    return torch.from_numpy(g.uniform(0, 1, n).astype(np.float64))


def _worker(rank, world, port, steps, spp, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    total = torch.zeros(64, dtype=torch.float64)
    seen = []
    for step in range(steps):
        begin, count = multigpu.sample_range(step, rank, world, spp)
        sums = torch.zeros(64, dtype=torch.float64)
        for s in range(begin, begin + count):
            sums += _fake_sample_image(s)
            seen.append(s)
        multigpu.reduce_sums([sums], dst=0)
        if rank == 0:
            total += sums
    ms = multigpu.max_over_ranks(10.0 + rank, torch.device("cpu"))
    gathered = [None] * world
    dist.all_gather_object(gathered, seen)
    if rank == 0:
        out.put((total.numpy(), sorted(sum(gathered, [])), ms))
    dist.destroy_process_group()


def test_two_rank_sample_sharding_and_reduce():
    world, steps, spp = 2, 3, 4
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, steps, spp, q)) for r in range(world)]
    for p in procs:
        p.start()
    total, seen, ms = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n = world * steps * spp
    assert seen == list(range(n))                       # every sample index exactly once
    want = sum(_fake_sample_image(s) for s in range(n)).numpy()
    assert np.allclose(total, want)
    assert np.allclose(multigpu.resolve(torch.from_numpy(total), n).numpy(), want / n)
    assert ms == 11.0                                   # max over ranks


def test_texel_ranges_partition():
    for count in (0, 1, 7, 1024 * 1024 + 3):
        for world in (1, 2, 8):
            pos = 0
            for r in range(world):
                b, n = multigpu.texel_range(count, r, world)
                assert b == pos
                pos += n
            assert pos == count


def _texel_worker(rank, world, port, covered, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # bench.py's C4 sharding: contiguous ranges of the covered texels, padded slabs gathered on rank 0
    begin, count = multigpu.texel_range(covered, rank, world)
    m = (covered + world - 1) // world
    slab = torch.zeros((m, 2), dtype=torch.float64)
    slab[:count, 0] = torch.arange(begin, begin + count, dtype=torch.float64)        # "baked" value = the texel's own index
    slab[:count, 1] = 1.0
    parts = [torch.zeros_like(slab) for _ in range(world)] if rank == 0 else None
    dist.gather(slab, parts, dst=0)
    if rank == 0:
        page = torch.full((covered,), -1.0, dtype=torch.float64)
        for r, part in enumerate(parts):
            b, c = multigpu.texel_range(covered, r, world)
            page[b:b + c] = part[:c, 0]
            assert float(part[:c, 1].sum()) == c
        out.put(page.numpy())
    dist.destroy_process_group()


def test_two_rank_texel_ranges_gather_to_one_page():
    world, covered = 2, 54437
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_texel_worker, args=(r, world, port, covered, q)) for r in range(world)]
    for p in procs:
        p.start()
    page = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.array_equal(page, np.arange(covered, dtype=np.float64))      # every covered texel exactly once, in place
