"""Multi-GPU frame assembly: pixels are independent (engine.rs:201-243 shares only an
immutable Arc<RayTracer>), so the frame is sharded by interleaved tile rows — tile row j (tile_rows
scan-lines) belongs to rank j % world — with the scene replicated on every GPU and NO collective on the
data path.  One exchange follows the render: an all-gather of each rank's compact u32 rows (NCCL over
NVLink on GPUs, gloo in the CPU tests), then a de-interleave into the reference's frame layout.

One process per GPU; torch.distributed is plumbing only.
"""
import ctypes as C

import numpy as np

from . import abi

DEFAULT_TILE_ROWS = 8


def owned_rows(height, tile_rows, world, rank):
    """Buffer rows (ascending) that `rank` owns; compact row k of the shard is owned_rows(...)[k].
    Mirrors rrt_shard_rows / buffer_row() in csrc (pure Python: usable without the CUDA library)."""
    rows = []
    j = rank
    while j * tile_rows < height:
        rows.extend(range(j * tile_rows, min(height, (j + 1) * tile_rows)))
        j += world
    return rows


def rows_per_block(height, tile_rows, world):
    """Padded row count of one gathered block = the largest shard (rank 0 owns the most)."""
    return max(len(owned_rows(height, tile_rows, world, r)) for r in range(world))


def deinterleave_host(gathered, width, height, tile_rows, world):
    """CPU statement of rrt_deinterleave_device: gathered = (world, rows_per_block, width) u32."""
    fb = np.zeros((height, width), dtype=np.uint32)
    for s in range(world):
        rows = owned_rows(height, tile_rows, world, s)
        fb[rows] = gathered[s, :len(rows)]
    fb[0] = 0     # buffer row 0 is never written by the reference (engine.rs:146-158)
    return fb


class ShardedRenderer:
    """Per-rank renderer: owns a replicated device scene, renders its interleaved tile rows, and
    assembles the full frame on every rank with one all-gather."""

    def __init__(self, ray_tracer, width, height, rank, world, tile_rows=DEFAULT_TILE_ROWS, group=None):
        import torch
        self.rt, self.W, self.H = ray_tracer, width, height
        self.rank, self.world, self.T = rank, world, tile_rows
        self.group = group
        self.shard = abi.Shard(tile_rows, world, rank)
        self.n_local = int(abi.lib().rrt_shard_rows(height, C.byref(self.shard)))
        self.rpb = rows_per_block(height, tile_rows, world)
        dev = torch.device("cuda", ray_tracer.device)
        self.device = dev
        self.local = torch.zeros((self.rpb, width), dtype=torch.int32, device=dev)     # u32 payload
        self.gathered = torch.zeros((world, self.rpb, width), dtype=torch.int32, device=dev) if world > 1 else None
        self.frame = torch.zeros((height, width), dtype=torch.int32, device=dev)

    def render(self, params, stream=None):
        """Enqueue this rank's share of Scene::draw_scene on `stream` (torch stream or None)."""
        import torch
        s = stream or torch.cuda.current_stream(self.device)
        out = self.local if self.world > 1 else self.frame
        abi.check(abi.lib().rrt_render_device(self.rt.handle, C.byref(params), C.byref(self.shard),
                                              out.data_ptr(), None, C.c_void_p(s.cuda_stream)))

    def assemble(self, stream=None):
        """all-gather the compact shards (NCCL) and scatter them into the reference frame layout."""
        import torch
        import torch.distributed as dist
        if self.world == 1:
            return self.frame
        dist.all_gather_into_tensor(self.gathered, self.local, group=self.group)
        s = stream or torch.cuda.current_stream(self.device)
        abi.check(abi.lib().rrt_deinterleave_device(self.rt.device, self.gathered.data_ptr(), self.rpb, self.W,
                                                    self.H, self.T, self.world, self.frame.data_ptr(),
                                                    C.c_void_p(s.cuda_stream)))
        return self.frame


def gather_frame_cpu(local_rows, width, height, tile_rows, rank, world, group=None):
    """gloo path used by the CPU tests: local_rows = (n_local, width) u32 numpy -> full frame."""
    import torch
    import torch.distributed as dist
    rpb = rows_per_block(height, tile_rows, world)
    pad = np.zeros((rpb, width), dtype=np.uint32)
    pad[:local_rows.shape[0]] = local_rows
    mine = torch.from_numpy(pad.view(np.int32))
    out = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    g = np.stack([o.numpy().view(np.uint32) for o in out])
    return deinterleave_host(g, width, height, tile_rows, world)
