"""The N > 1 path on CPU: two ranks over gloo run the bench's world partition + statistics
reduction (the only collective in the path), each stepping ITS worlds with the CPU oracle, and
the gathered totals must equal a single process stepping all worlds."""
import os
import socket
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NW_PER_RANK, BPW, FRAMES = 3, 96, 45


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _step_worlds(first, count):
    import starframe_b200 as sf
    from oracle import loader
    from starframe_b200 import scenes

    bodies = scenes.worlds(count, BPW, first_world=first)
    pairs = contacts = 0.0
    checksum = 0.0
    for k in range(count):
        w = sf.World(loader.api())
        w.upload_bodies(bodies[k])
        for _ in range(FRAMES):
            w.step()
        st = w.stats()
        pairs += st["n_pairs"]
        contacts += st["n_contacts"]
        s = w.download_bodies()
        checksum += float(np.sum(s["x"].astype(np.float64)) + np.sum(s["y"].astype(np.float64)))
    return pairs, contacts, checksum


def _rank_main(rank, world_size, port, out):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    import bench

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank,
                            world_size=world_size)
    mine = bench.world_range(rank, world_size, world_size * NW_PER_RANK)
    pairs, contacts, checksum = _step_worlds(mine.start, len(mine))
    bench.barrier(world_size)
    tot = bench.allreduce_sum([pairs, contacts, checksum, float(len(mine))], world_size)
    slowest = bench.allreduce_max(float(rank + 1), world_size)
    if rank == 0:
        out.put((tot, slowest, list(mine)))
    dist.destroy_process_group()


def test_two_ranks_partition_and_reduce_like_one_process():
    sys.path.insert(0, ROOT)
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    tot, slowest, rank0_worlds = out.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    pairs, contacts, checksum = _step_worlds(0, 2 * NW_PER_RANK)
    assert rank0_worlds == [0, 1, 2]
    assert tot[3] == 2 * NW_PER_RANK
    assert tot[0] == pairs and tot[1] == contacts
    assert abs(tot[2] - checksum) < 1e-6 * abs(checksum)
    assert slowest == 2.0  # MAX over ranks, the rule for every multi-GPU time


def test_world_ranges_tile_the_world_list():
    sys.path.insert(0, ROOT)
    import bench

    for ranks in (1, 2, 3, 4, 8):
        seen = []
        for r in range(ranks):
            seen += list(bench.world_range(r, ranks, 8192))
        assert seen == list(range(8192))  # BASELINE.json config 5: 8192 worlds over 1/2/4/8 GPUs
        sizes = [len(bench.world_range(r, ranks, 8192)) for r in range(ranks)]
        assert max(sizes) - min(sizes) <= 1
