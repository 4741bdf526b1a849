"""Job sharding across ranks (world_size 2, gloo, CPU): the N>1 host logic of bench.py / BASELINE configs[4]."""
import os
import socket

import torch.multiprocessing as mp

from imagestitching_b200.shard import shard_jobs


def test_shard_partition():
    for n in (0, 1, 7, 10, 10000):
        for world in (1, 2, 4, 8):
            seen = []
            for r in range(world):
                part = shard_jobs(n, r, world)
                seen += part
                assert len(part) in (n // world, n // world + 1)
            assert sorted(seen) == list(range(n))


def _worker(rank, world, port, q):
    import torch.distributed as dist

    from imagestitching_b200 import compute_layout
    from imagestitching_b200.shard import digest, gather_digests, shard_jobs

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_jobs = 11
    local = {}
    for j in shard_jobs(n_jobs, rank, world):
        # the per-job host work every rank does on its own: the layout twin (pixels need a GPU)
        dims = [(4032 - 16 * j, 3024)] * 9
        lay = compute_layout([{"naturalWidth": w, "naturalHeight": h} for (w, h) in dims], "vertical")
        local[j] = digest(repr((lay["canvasW"], lay["canvasH"], lay["rects"])).encode())
    merged = gather_digests(local, world)
    dist.barrier()
    if rank == 0:
        q.put(merged)
    dist.destroy_process_group()


def test_two_ranks_cover_all_jobs_once():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    merged = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(merged) == list(range(11))
    # same digests as a single-rank run
    from imagestitching_b200 import compute_layout
    from imagestitching_b200.shard import digest
    for j in range(11):
        dims = [(4032 - 16 * j, 3024)] * 9
        lay = compute_layout([{"naturalWidth": w, "naturalHeight": h} for (w, h) in dims], "vertical")
        assert merged[j] == digest(repr((lay["canvasW"], lay["canvasH"], lay["rects"])).encode())
