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


class HostFrame:
    """One canvas in host memory shared by the ranks of a node (include/rrt.h, rrt_host_frame_*): a POSIX
    shared-memory segment every rank maps and page-locks, with a process-shared barrier in its header."""

    def __init__(self, name, width, height, n_ranks, create, device=None, timeout=60.0):
        L = abi.lib()
        h = C.c_void_p()
        if create:
            abi.check(L.rrt_host_frame_create(name.encode(), width, height, n_ranks, C.byref(h)))
        else:
            abi.check(L.rrt_host_frame_open(name.encode(), float(timeout), C.byref(h)))
        self._h, self.name, self.W, self.H = h, name, width, height
        if device is not None:
            abi.check(L.rrt_host_frame_pin(h, device))
        self.ptr = L.rrt_host_frame_pixels(h)
        buf = (C.c_uint32 * (width * height)).from_address(self.ptr)
        self.pixels = np.frombuffer(buf, dtype=np.uint32).reshape(height, width)

    def barrier(self, timeout=120.0):
        abi.check(abi.lib().rrt_host_frame_barrier(self._h, float(timeout)))

    def close(self):
        if self._h:
            self.pixels = None
            abi.lib().rrt_host_frame_close(self._h)
            self._h = None


class ShardedRenderer:
    """Per-rank renderer: owns a replicated device scene and renders its interleaved tile rows.

    assemble="nccl": compact rows, one all-gather + de-interleave, the full frame on EVERY rank's device.
    assemble="peer": fused render + gather — rank 0 owns the frame(s) (rrt_frame_alloc), the other ranks map them
        through CUDA IPC and their render kernels store finished pixels straight into rank 0's memory over NVLink;
        the only collective left is a 4-byte all-reduce, enqueued on the render stream, that orders "all kernels
        of this frame done".  TWO frames are used alternately (frame k lands in buffer k & 1): the token of frame
        k+1 can only complete after rank 0 enqueued it, i.e. after whatever rank 0 put on the stream to consume
        frame k, so by the time any rank stores frame k+2 into the buffer of frame k that consumer has run.
        Rule for the caller: consume `frame` on the render stream (or make that stream wait for the consumer)
        before calling render() again.
    assemble="host": the frame is wanted in HOST memory (the reference's Canvas.buffer, engine.rs:123-128), in one
        page-locked shared-memory frame (HostFrame) that every rank maps.  Default: every rank's render kernel stores
        the pixels of ITS rows straight into that frame over its own PCIe link (rrt_render_to_host_frame); render() first
        waits for "rank 0 is done reading the previous frame", assemble() = a stream sync and the barrier "frame
        complete".  With RRT_HOST_STORES=0: render compactly into device memory, then (behind the first barrier, which
        then hides behind the kernel) copy the rows out.  No device-side exchange at all either way.
    """

    def __init__(self, ray_tracer, width, height, rank, world, tile_rows=DEFAULT_TILE_ROWS, group=None,
                 assemble="nccl"):
        import torch
        self.rt, self.W, self.H = ray_tracer, width, height
        self.rank, self.world, self.T = rank, world, tile_rows
        self.group = group
        self.mode = assemble if (world > 1 or assemble == "host") else "nccl"
        self.shard = abi.Shard(tile_rows, world, rank)
        self.n_local = int(abi.lib().rrt_shard_rows(height, C.byref(self.shard)))
        self.rpb = rows_per_block(height, tile_rows, world)
        dev = torch.device("cuda", ray_tracer.device)
        self.device = dev
        self._own_frames, self._peer_frames, self._host = [], [], None
        self._k = 0                 # frames rendered so far (peer: selects the buffer)
        self._outstanding = False   # host: a completed frame has not been released yet
        import os
        self._direct = os.environ.get("RRT_HOST_STORES", "1") != "0"   # host: kernels store into the canvas themselves
        # set-up, not first-frame work: the accumulators and the dispatch-order arrays of this shape (rrt_scene_reserve)
        p0 = ray_tracer.params(width, height)
        abi.check(abi.lib().rrt_scene_reserve(ray_tracer.handle, C.byref(p0), C.byref(self.shard),
                                              1 if (self.mode == "host" and not self._direct) else 0))
        if self.mode == "peer":
            self._setup_peer_frames()
            return
        if self.mode == "host":
            self._setup_host_frame()
            return
        self.local = torch.zeros((self.rpb, width), dtype=torch.int32, device=dev)     # u32 payload
        self.gathered = torch.zeros((world, self.rpb, width), dtype=torch.int32, device=dev) if world > 1 else None
        self.frame = torch.zeros((height, width), dtype=torch.int32, device=dev)

    def _setup_host_frame(self):
        import os
        import torch.distributed as dist
        box = [f"/rrt_frame_{os.getpid()}_{id(self) & 0xFFFFFF:x}" if self.rank == 0 else None]
        if self.rank == 0:
            self._host = HostFrame(box[0], self.W, self.H, self.world, create=True, device=self.rt.device)
        if self.world > 1:
            dist.broadcast_object_list(box, src=0, group=self.group)
        if self.rank != 0:
            self._host = HostFrame(box[0], self.W, self.H, self.world, create=False, device=self.rt.device)
        self.frame = self._host.pixels        # numpy (H, W) u32 view of the shared canvas
        import torch
        self._rows = torch.zeros((max(self.n_local, 1), self.W), dtype=torch.int32, device=self.device)   # this rank's compact rows

    def _setup_peer_frames(self):
        import torch
        import torch.distributed as dist
        L = abi.lib()
        handles = []
        if self.rank == 0:
            for _ in range(2):
                p = C.c_void_p()
                abi.check(L.rrt_frame_alloc(self.rt.device, self.W, self.H, C.byref(p)))
                self._own_frames.append(p.value)
                handle = (C.c_ubyte * 64)()
                abi.check(L.rrt_ipc_get_handle(self.rt.device, p, handle))
                handles.append(bytes(handle))
        box = [handles]
        dist.broadcast_object_list(box, src=0, group=self.group)
        if self.rank == 0:
            self._dest = list(self._own_frames)
            self._frames = [torch.as_tensor(_DevFrame(d, self.H, self.W), device=self.device) for d in self._dest]
        else:
            for hb in box[0]:
                q = C.c_void_p()
                buf = (C.c_ubyte * 64).from_buffer_copy(hb)
                abi.check(L.rrt_ipc_open_handle(self.rt.device, buf, C.byref(q)))
                self._peer_frames.append(q.value)
            self._dest = list(self._peer_frames)
            self._frames = [None, None]
        self.frame = self._frames[0]
        self._token = torch.zeros(1, dtype=torch.int32, device=self.device)

    def close(self):
        L = abi.lib()
        for q in self._peer_frames:
            L.rrt_ipc_close_handle(self.rt.device, C.c_void_p(q))
        self._peer_frames = []
        if self._own_frames:
            self.frame = None
            self._frames = [None, None]
            for p in self._own_frames:
                L.rrt_frame_free(self.rt.device, C.c_void_p(p))
            self._own_frames = []
        if self._host is not None:
            self.frame = None
            self._host.close()
            self._host = None

    def render(self, params, stream=None):
        """Enqueue this rank's share of Scene::draw_scene on `stream` (torch stream or None = current)."""
        import torch
        s = stream or torch.cuda.current_stream(self.device)
        if self.mode == "peer":
            self._cur = self._k & 1
            self._k += 1
            abi.check(abi.lib().rrt_render_device_into_frame(self.rt.handle, C.byref(params), C.byref(self.shard),
                                                             C.c_void_p(self._dest[self._cur]), C.c_void_p(s.cuda_stream)))
            return
        if self.mode == "host":
            if self._direct:
                # the kernel stores its pixels straight into the shared page-locked canvas (rrt_render_to_host_frame,
                # direct host stores): it may only start once rank 0 has finished reading the previous frame
                if self._outstanding and self.world > 1:
                    self._host.barrier()
                abi.check(abi.lib().rrt_render_to_host_frame(self.rt.handle, C.byref(params), C.byref(self.shard),
                                                             C.c_void_p(self._host.ptr), C.c_void_p(s.cuda_stream)))
                return
            # render first; "rank 0 has finished reading the previous frame" is only needed before the COPY (assemble)
            abi.check(abi.lib().rrt_render_device(self.rt.handle, C.byref(params), C.byref(self.shard),
                                                  self._rows.data_ptr(), None, C.c_void_p(s.cuda_stream)))
            return
        out = self.local if self.world > 1 else self.frame
        abi.check(abi.lib().rrt_render_device(self.rt.handle, C.byref(params), C.byref(self.shard),
                                              out.data_ptr(), None, C.c_void_p(s.cuda_stream)))

    def assemble(self, stream=None):
        """nccl: all-gather the compact shards and scatter them into the reference frame layout.
        peer: the pixels are already in rank 0's frame; order every rank's kernel before what follows on `stream`.
        host: wait for this rank's copies, then for every rank's ("frame complete"); returns the shared host frame."""
        import torch
        import torch.distributed as dist
        s = stream or torch.cuda.current_stream(self.device)
        if self.mode == "host":
            if not self._direct:
                if self._outstanding and self.world > 1:
                    self._host.barrier()      # rank 0 has finished reading the previous frame (waited for behind the kernel)
                abi.check(abi.lib().rrt_copy_shard_to_host_frame(self.rt.device, self._rows.data_ptr(), self.W, self.H,
                                                                 C.byref(self.shard), C.c_void_p(self._host.ptr),
                                                                 C.c_void_p(s.cuda_stream)))
            s.synchronize()
            if self.world > 1:
                self._host.barrier()          # every rank's rows have landed: the frame is complete
            self._outstanding = True
            return self.frame
        if self.world == 1:
            return self.frame
        if self.mode == "peer":
            with torch.cuda.stream(s):
                dist.all_reduce(self._token, group=self.group)     # stream-ordered after each rank's render kernel
            self.frame = self._frames[self._cur]
            return self.frame
        with torch.cuda.stream(s):
            dist.all_gather_into_tensor(self.gathered, self.local, group=self.group)
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


# ---------------------------------------------------------------- frame-parallel sequences (SURVEY §8e "Orbit", §8 f4)
# The frames of an animation are as independent as the pixels of a frame (the scene is immutable, only the camera
# moves): frame k of a sequence belongs to rank k % world, each rank renders WHOLE frames with the multi-frame call
# (rrt_render_frames: frame i leaves for the host while frame i+1 renders) and nothing is exchanged until the
# sequence is collected.  No per-frame gather, no per-frame drain tail; the price is latency (a frame is as slow as
# one GPU makes it), so this is the mode for offline sequences such as config 5's 64-frame orbit, tile rows the mode
# for one frame as fast as possible.

def owned_frames(n_frames, world, rank):
    """Indices (ascending) of the frames `rank` renders: k % world == rank."""
    return list(range(rank, n_frames, world))


def gather_frames(local, n_frames, rank, world, dst=0, group=None, device=None):
    """Collect a frame-parallel sequence on rank `dst`: local = (len(owned_frames), H, W) u32 numpy array of this
    rank's frames in ascending frame order -> (n_frames, H, W) on dst, None elsewhere.  One gather of equally sized
    (padded) blocks; NCCL needs device tensors (`device`), gloo takes the host arrays as they are."""
    import torch
    import torch.distributed as dist
    mine = owned_frames(n_frames, world, rank)
    if local.dtype != np.uint32 or local.ndim != 3 or local.shape[0] != len(mine):
        raise ValueError("local must be (len(owned_frames), H, W) uint32")
    H, W = local.shape[1:]
    per = (n_frames + world - 1) // world            # rank 0 owns the most
    pad = np.zeros((per, H, W), dtype=np.uint32)
    pad[:len(mine)] = local
    t = torch.from_numpy(pad.view(np.int32))
    if device is not None:
        t = t.to(device)
    if world == 1:
        blocks = [t]
    else:
        blocks = [torch.empty_like(t) for _ in range(world)] if rank == dst else None
        dist.gather(t, blocks, dst=dst, group=group)
    if rank != dst:
        return None
    out = np.zeros((n_frames, H, W), dtype=np.uint32)
    for r in range(world):
        fr = owned_frames(n_frames, world, r)
        out[fr] = blocks[r].cpu().numpy().view(np.uint32)[:len(fr)]
    return out


class FrameParallelSequence:
    """One process per GPU, scene replicated (as ShardedRenderer): render_local() renders this rank's frames of a
    sequence through Scene.draw_frames, collect() gathers the sequence on one rank."""

    def __init__(self, ray_tracer, width, height, rank, world, group=None):
        self.rt, self.W, self.H = ray_tracer, width, height
        self.rank, self.world, self.group = rank, world, group
        self.device_ms = None

    def render_local(self, overrides_per_frame, out=None):
        """overrides_per_frame: the WHOLE sequence (every rank passes the same list); -> this rank's frames,
        (len(owned_frames), H, W) u32, frame owned_frames(...)[i] at [i]."""
        from .scene import Scene
        mine = owned_frames(len(overrides_per_frame), self.world, self.rank)
        sc = Scene(self.W, self.H)
        frames = sc.draw_frames(self.rt, [overrides_per_frame[k] for k in mine], out=out)
        self.device_ms = sc.device_ms
        return frames

    def collect(self, local, n_frames, dst=0):
        import torch
        import torch.distributed as dist
        device = None
        if self.world > 1 and dist.get_backend(self.group) == "nccl":
            device = torch.device("cuda", self.rt.device)
        return gather_frames(local, n_frames, self.rank, self.world, dst=dst, group=self.group, device=device)
