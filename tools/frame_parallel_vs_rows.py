"""Sequences on N GPUs, two ways, measured on hardware:
  rows   — every frame sharded by interleaved tile rows over the ranks, assembled in one shared host canvas
           (ShardedRenderer(assemble="host")): lowest latency per frame;
  frames — frame k rendered whole by rank k % N through rrt_render_frames (FrameParallelSequence), the sequence
           gathered on rank 0: no per-frame exchange at all.
Both must give the same frames (blake2b of every frame, compared on rank 0).
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
      tools/frame_parallel_vs_rows.py [c4] [n_frames]
"""
import hashlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402
import rrt_b200  # noqa: E402
from rrt_b200.sharding import DEFAULT_TILE_ROWS, FrameParallelSequence, ShardedRenderer  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c4"
n_frames = int(sys.argv[2]) if len(sys.argv) > 2 else 32
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
bench.set_openmp_threads(bench.host_threads() // world)
desc, W, H = bench.WORKLOADS[name]
rt = rrt_b200.RayTracer(bench.load_workload(name), device=local)
cams = []
for k in range(n_frames):
    origin, rot = rrt_b200.orbit_camera(k % 64)
    cams.append(dict(origin=origin, camera_rot=rot))


def digest(a):
    return hashlib.blake2b(np.ascontiguousarray(a, dtype="<u4").tobytes(), digest_size=8).hexdigest()


def barrier():
    dist.barrier()
    torch.cuda.synchronize(dev)


# ---- rows: every frame over all ranks
sr = ShardedRenderer(rt, W, H, rank, world, tile_rows=DEFAULT_TILE_ROWS, assemble="host")
rows_digests, lat = [], []
for phase in ("warm", "timed", "verify"):
    barrier()
    t0 = time.perf_counter()
    for c in cams:
        t1 = time.perf_counter()
        sr.render(rt.params(W, H, **c))
        frame = sr.assemble()
        if phase == "timed":
            lat.append(time.perf_counter() - t1)
        if phase == "verify" and rank == 0:
            rows_digests.append(digest(frame))       # rank 0 reads the canvas before it enters the next frame
    barrier()
    if phase == "timed":
        rows_s = time.perf_counter() - t0
sr.close()

# ---- frames: frame k on rank k % world
fps = FrameParallelSequence(rt, W, H, rank, world)
mine = len(range(rank, n_frames, world))
pinned = torch.empty((max(mine, 1), H, W), dtype=torch.int32, pin_memory=True)
out = pinned.numpy().view(np.uint32)[:mine]
for warm in (True, False):
    barrier()
    t0 = time.perf_counter()
    local_frames = fps.render_local(cams, out=out)
    t_render = time.perf_counter() - t0
    seq = fps.collect(local_frames, n_frames)
    barrier()
    frames_s = time.perf_counter() - t0
t = torch.tensor([rows_s, frames_s, t_render], dtype=torch.float64, device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    frames_digests = [digest(seq[k]) for k in range(n_frames)]
    print(json.dumps({
        "workload": f"{name}: {desc} at {W}x{H}, {n_frames} frames of the orbit", "n_gpus": world,
        "rows": {"seconds": float(t[0]), "frames_per_s": n_frames / float(t[0]), "Mrays_per_s": 4e-6 * W * H * n_frames / float(t[0]),
                 "latency_ms_per_frame": 1e3 * sum(lat) / len(lat)},
        "frames": {"seconds": float(t[1]), "frames_per_s": n_frames / float(t[1]), "Mrays_per_s": 4e-6 * W * H * n_frames / float(t[1]),
                   "render_seconds_slowest_rank": float(t[2]), "latency_ms_per_frame": 1e3 * float(t[2]) / max(mine, 1),
                   "note": "seconds include the gather of the whole sequence on rank 0"},
        "digests_equal": rows_digests == frames_digests, "first_digest": rows_digests[0]}), flush=True)
rt.close()
dist.destroy_process_group()
