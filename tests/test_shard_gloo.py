"""N > 1 host logic on CPU: two gloo ranks shard a stream of frame pairs (orb_slam_tracking_b200.shard), each
runs its block through the CPU oracle (standing in for the per-rank CUDA context: no GPU here), rank 0 gathers
and the union must equal the single-process result; max-over-ranks timing reduction."""
import os
import socket

import numpy as np
import pytest

from orb_slam_tracking_b200 import shard


def test_shard_pairs_partition():
    for n in (0, 1, 7, 8, 256, 1001):
        for world in (1, 2, 3, 4, 8):
            blocks = [shard.shard_pairs(n, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            for a, b in zip(blocks, blocks[1:]):
                assert a[1] == b[0]
            sizes = [hi - lo for lo, hi in blocks]
            assert max(sizes) - min(sizes) <= 1
            assert shard.shard_frames(n, world, 0)[0] == 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_pairs, W, H, nf, q):
    import torch.distributed as dist
    from oracle import orb_oracle as O
    from orb_slam_tracking_b200 import synth
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard.shard_pairs(n_pairs, world, rank)
    K4, D4 = synth.camera_for(W, H)
    bounds = O.image_bounds(W, H, K4, D4)
    E = O.Extractor(nf, 1.2, 8, 20, 7)
    mine = []
    for p in range(lo, hi):
        F = []
        for im in synth.frame_pair(p, W, H):
            k, d, _ = E(im)
            F.append(O.Frame(k, d, K4, D4, bounds))
        n, m12, _ = O.search_for_initialization(F[0], F[1], 0.9, True, 100)
        mine.append((p, n, m12.tobytes()))
    dist.barrier()
    t = shard.max_over_ranks(10.0 + rank)      # rank 1 is "slower": the reduction must return its time
    allr = shard.gather_to_rank0(mine)
    if rank == 0:
        q.put((t, [x for part in allr for x in part]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sharded_stream_matches_single_process(oracle):
    import torch.multiprocessing as mp
    from orb_slam_tracking_b200 import synth
    n_pairs, W, H, nf, world = 5, 320, 240, 400, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_pairs, W, H, nf, q)) for r in range(world)]
    for p in procs:
        p.start()
    t, got = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert t == 11.0
    assert [g[0] for g in got] == list(range(n_pairs))  # contiguous blocks, rank order
    K4, D4 = synth.camera_for(W, H)
    bounds = oracle.image_bounds(W, H, K4, D4)
    E = oracle.Extractor(nf, 1.2, 8, 20, 7)
    for p, n, m in got:
        F = []
        for im in synth.frame_pair(p, W, H):
            k, d, _ = E(im)
            F.append(oracle.Frame(k, d, K4, D4, bounds))
        rn, rm12, _ = oracle.search_for_initialization(F[0], F[1], 0.9, True, 100)
        assert n == rn and m == rm12.tobytes(), p
