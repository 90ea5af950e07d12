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


class _DevFrame:
    """A raw device frame (rrt_frame_alloc / rrt_ipc_open_handle) exposed to torch without a copy."""

    def __init__(self, ptr, height, width):
        self.ptr = ptr
        self.__cuda_array_interface__ = {"shape": (height, width), "typestr": "<i4", "data": (ptr, False), "version": 2}


class ShardedRenderer:
    """Per-rank renderer: owns a replicated device scene and renders its interleaved tile rows.

    assemble="nccl": compact rows, one all-gather + de-interleave, the full frame on EVERY rank.
    assemble="peer": fused render + gather — rank 0 owns the frame (rrt_frame_alloc), the other ranks map it
        through CUDA IPC and their render kernels store finished pixels straight into it over NVLink; the
        only collective left is a 4-byte all-reduce that orders "all kernels done" on the streams.
    """

    def __init__(self, ray_tracer, width, height, rank, world, tile_rows=DEFAULT_TILE_ROWS, group=None,
                 assemble="nccl"):
        import torch
        self.rt, self.W, self.H = ray_tracer, width, height
        self.rank, self.world, self.T = rank, world, tile_rows
        self.group = group
        self.mode = assemble if world > 1 else "nccl"
        self.shard = abi.Shard(tile_rows, world, rank)
        self.n_local = int(abi.lib().rrt_shard_rows(height, C.byref(self.shard)))
        self.rpb = rows_per_block(height, tile_rows, world)
        dev = torch.device("cuda", ray_tracer.device)
        self.device = dev
        self._own_frame = self._peer_frame = None
        if self.mode == "peer":
            self._setup_peer_frame()
            return
        self.local = torch.zeros((self.rpb, width), dtype=torch.int32, device=dev)     # u32 payload
        self.gathered = torch.zeros((world, self.rpb, width), dtype=torch.int32, device=dev) if world > 1 else None
        self.frame = torch.zeros((height, width), dtype=torch.int32, device=dev)

    def _setup_peer_frame(self):
        import torch
        import torch.distributed as dist
        L = abi.lib()
        handle = (C.c_ubyte * 64)()
        if self.rank == 0:
            p = C.c_void_p()
            abi.check(L.rrt_frame_alloc(self.rt.device, self.W, self.H, C.byref(p)))
            self._own_frame = p.value
            abi.check(L.rrt_ipc_get_handle(self.rt.device, p, handle))
        box = [bytes(handle)]
        dist.broadcast_object_list(box, src=0, group=self.group)
        if self.rank == 0:
            self.dest = self._own_frame
            self.frame = torch.as_tensor(_DevFrame(self.dest, self.H, self.W), device=self.device)
        else:
            q = C.c_void_p()
            buf = (C.c_ubyte * 64).from_buffer_copy(box[0])
            abi.check(L.rrt_ipc_open_handle(self.rt.device, buf, C.byref(q)))
            self._peer_frame = q.value
            self.dest = q.value
            self.frame = None
        self._token = torch.zeros(1, dtype=torch.int32, device=self.device)

    def close(self):
        L = abi.lib()
        if self._peer_frame:
            L.rrt_ipc_close_handle(self.rt.device, C.c_void_p(self._peer_frame))
            self._peer_frame = None
        if self._own_frame:
            self.frame = None
            L.rrt_frame_free(self.rt.device, C.c_void_p(self._own_frame))
            self._own_frame = None

    def render(self, params, stream=None):
        """Enqueue this rank's share of Scene::draw_scene on `stream` (torch stream or None)."""
        import torch
        s = stream or torch.cuda.current_stream(self.device)
        if self.mode == "peer":
            abi.check(abi.lib().rrt_render_device_into_frame(self.rt.handle, C.byref(params), C.byref(self.shard),
                                                             C.c_void_p(self.dest), C.c_void_p(s.cuda_stream)))
            return
        out = self.local if self.world > 1 else self.frame
        abi.check(abi.lib().rrt_render_device(self.rt.handle, C.byref(params), C.byref(self.shard),
                                              out.data_ptr(), None, C.c_void_p(s.cuda_stream)))

    def assemble(self, stream=None):
        """nccl: all-gather the compact shards and scatter them into the reference frame layout.
        peer: the pixels are already in rank 0's frame; order every rank's kernel before what follows."""
        import torch
        import torch.distributed as dist
        if self.world == 1:
            return self.frame
        if self.mode == "peer":
            dist.all_reduce(self._token, group=self.group)     # stream-ordered after each rank's render kernel
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
