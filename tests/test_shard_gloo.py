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


# ---- one pair's Hamming matrix sharded by query rows (brute-force matching, BASELINE.json configs[4]) ----------------
def test_shard_rows_partition():
    for n in (0, 1, 433, 1737):
        for world in (1, 2, 8):
            blocks = [shard.shard_rows(n, world, r) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n and all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))


def _pair_frames(O, W, H, nf, pair):
    from orb_slam_tracking_b200 import synth
    K4, D4 = synth.camera_for(W, H)
    bounds = O.image_bounds(W, H, K4, D4)
    E = O.Extractor(nf, 1.2, 8, 20, 7)
    F = []
    for im in synth.frame_pair(pair, W, H):
        k, d, _ = E(im)
        F.append(O.Frame(k, d, K4, D4, bounds))
    return F


def _row_worker(rank, world, port, W, H, nf, window, q):
    import sys
    import torch.distributed as dist
    from oracle import orb_oracle as O
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import bf_ref
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    F = _pair_frames(O, W, H, nf, 3)                       # every rank holds both frames (F2 replicated, no collective)
    q0, q1 = shard.shard_rows(len(bf_ref.level0(F[0])), world, rank)
    mine = bf_ref.rows(F[0], F[1], window, q0, q1)         # stands in for orbx_match_rows on this rank's GPU
    parts = shard.gather_to_rank0((q0, mine.tobytes()))
    if rank == 0:
        table = np.concatenate([np.frombuffer(b, np.uint32).reshape(-1, 8) for _, b in sorted(parts)])
        n, m12, cnt = bf_ref.resolve(F[0], F[1], table, 0.9, True, window)   # stands in for orbx_match_resolve
        q.put((n, m12.tobytes(), cnt.tobytes()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("window", [8192, 60])
def test_two_rank_row_sharded_match_equals_the_reference_loop(oracle, window):
    """The top-list + ordered-resolve scheme (csrc/match_bf.cu) over two gloo ranks == SearchForInitialization."""
    import torch.multiprocessing as mp
    W, H, nf, world = 320, 240, 500, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_row_worker, args=(r, world, port, W, H, nf, window, q)) for r in range(world)]
    for p in procs:
        p.start()
    n, m12, cnt = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    F = _pair_frames(oracle, W, H, nf, 3)
    rn, rm12, rcnt = oracle.search_for_initialization(F[0], F[1], 0.9, True, window)
    assert n == rn and m12 == rm12.tobytes() and cnt == rcnt.tobytes()
