"""N>1 host logic on CPU: world_size-2 (and 3) gloo groups assemble a frame from interleaved tile-row
shards exactly as the NCCL path does on GPUs (rust-ray-tracer_b200/sharding.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _pattern(rows, W):
    r = np.asarray(rows, dtype=np.uint32)[:, None]
    c = np.arange(W, dtype=np.uint32)[None, :]
    return (r * 4099 + c * 7 + 1) & 0x00FFFFFF


def _worker(rank, world, port, W, H, T, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import rrt_b200  # noqa: F401
    from rrt_b200.sharding import gather_frame_cpu, owned_rows
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    rows = owned_rows(H, T, world, rank)
    local = _pattern(rows, W)                      # stands in for this rank's rendered compact rows
    frame = gather_frame_cpu(local, W, H, T, rank, world)
    expect = _pattern(range(H), W)
    expect[0] = 0                                  # the reference never writes buffer row 0
    q.put((rank, bool(np.array_equal(frame, expect)), len(rows)))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,W,H,T", [(2, 64, 45, 8), (2, 16, 7, 3), (3, 32, 50, 4)])
def test_gloo_gather_deinterleave(world, W, H, T):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, world, port, W, H, T, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    assert sum(n for _, _, n in res) == H


def test_deinterleave_host_matches_shard_mapping():
    sys.path.insert(0, ROOT)
    import rrt_b200  # noqa: F401
    from rrt_b200.sharding import deinterleave_host, owned_rows, rows_per_block
    W, H, T, G = 8, 37, 4, 4
    rpb = rows_per_block(H, T, G)
    g = np.zeros((G, rpb, W), np.uint32)
    for s in range(G):
        rows = owned_rows(H, T, G, s)
        g[s, :len(rows)] = _pattern(rows, W)
    e = _pattern(range(H), W)
    e[0] = 0
    assert np.array_equal(deinterleave_host(g, W, H, T, G), e)


# ---------------------------------------------------------------- frame-parallel sequences
def _frame_pattern(k, W, H):
    return (_pattern(range(H), W) + np.uint32(k * 65537)) & 0x00FFFFFF


def _seq_worker(rank, world, port, n, W, H, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import rrt_b200  # noqa: F401
    from rrt_b200.sharding import gather_frames, owned_frames
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    mine = owned_frames(n, world, rank)
    local = np.stack([_frame_pattern(k, W, H) for k in mine]) if mine else np.zeros((0, H, W), np.uint32)
    seq = gather_frames(local, n, rank, world, dst=0)
    ok = seq is None if rank != 0 else all(np.array_equal(seq[k], _frame_pattern(k, W, H)) for k in range(n)) and seq.shape == (n, H, W)
    q.put((rank, bool(ok), len(mine)))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 5), (3, 7), (3, 2)])
def test_gloo_frame_parallel_sequence(world, n):
    # frame k is rendered by rank k % world; one gather of padded blocks puts the sequence together on rank 0 —
    # ragged shares (n not a multiple of world) and a rank without any frame included
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_seq_worker, args=(r, world, port, n, 24, 10, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    assert sum(m for _, _, m in res) == n


def test_owned_frames_partition():
    sys.path.insert(0, ROOT)
    import rrt_b200  # noqa: F401
    from rrt_b200.sharding import gather_frames, owned_frames
    for world in (1, 2, 3, 8):
        for n in (0, 1, 7, 64):
            parts = [owned_frames(n, world, r) for r in range(world)]
            assert sorted(k for p in parts for k in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    one = np.stack([_frame_pattern(k, 6, 4) for k in range(3)])
    assert np.array_equal(gather_frames(one, 3, 0, 1), one)       # world 1: no process group needed
