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


# ---------------------------------------------------------------- the shared host canvas (rrt_host_frame_*, no GPU needed)
def _host_frame_worker(rank, world, name, W, H, T, q):
    sys.path.insert(0, ROOT)
    import rrt_b200  # noqa: F401
    from rrt_b200 import abi
    from rrt_b200.sharding import HostFrame, ShardedRenderer, owned_rows  # noqa: F401  (import check of the N>1 classes)
    import ctypes as C
    hf = HostFrame(name, W, H, world, create=(rank == 0))           # no device: not page-locked, same protocol
    rows = owned_rows(H, T, world, rank)
    ok = True
    for frame in range(3):
        if frame:
            hf.barrier(30.0)                                         # "rank 0 is done reading the previous frame"
        hf.pixels[rows] = _pattern(rows, W) + frame                  # stands in for this rank's device->host row copies
        hf.barrier(30.0)                                             # "frame complete"
        if rank == 0:
            expect = _pattern(range(H), W) + frame
            ok = ok and bool(np.array_equal(hf.pixels, expect))
    # the CUDA IPC handle slot of the header: rank 0 publishes 64 bytes, the others fetch them
    L = abi.lib()
    buf = (C.c_ubyte * 64)()
    if rank == 0:
        for i in range(64):
            buf[i] = (i * 7 + 3) & 255
        abi.check(L.rrt_host_frame_publish_handle(hf._h, buf))
    else:
        abi.check(L.rrt_host_frame_fetch_handle(hf._h, 30.0, buf))
        ok = ok and bytes(buf) == bytes((i * 7 + 3) & 255 for i in range(64))
    hf.barrier(30.0)
    hf.close()
    q.put((rank, ok))


@pytest.mark.parametrize("world,W,H,T", [(2, 64, 45, 8), (3, 40, 50, 4)])
def test_shared_host_frame_barrier_and_handle_slot(world, W, H, T):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    name = f"/rrt_test_{os.getpid()}_{world}"
    ps = [ctx.Process(target=_host_frame_worker, args=(r, world, name, W, H, T, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok in res)
    assert not os.path.exists("/dev/shm" + name)                     # the creator unlinked the segment


def test_host_frame_barrier_times_out_instead_of_hanging(rrt):
    from rrt_b200.sharding import HostFrame
    hf = HostFrame(f"/rrt_test_timeout_{os.getpid()}", 8, 4, 2, create=True)    # 2 ranks expected, 1 present
    with pytest.raises(Exception) as ei:
        hf.barrier(0.3)
    assert "timeout" in str(ei.value)
    hf.close()
