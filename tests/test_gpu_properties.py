"""Size-independent properties at BASELINE.json's full sizes (the oracle would take minutes there)."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def render(rrt, rt, W, H, hits=False):
    sc = rrt.Scene(W, H)
    fb = sc.draw_scene(rt, want_hit_ids=hits).copy()
    return fb, sc.hit_ids


def test_config2_culled_equals_linear_scan(rrt, scenes, monkeypatch):
    # config[1]: model2 at 1920x1080. The culling hierarchies are a conservative filter in front of the
    # exact f64 test, so the frame must be IDENTICAL with them switched off.
    sd = scenes("model2")
    rt = rrt.RayTracer(sd)
    fb, hit = render(rrt, rt, 1920, 1080, hits=True)
    fb2, _ = render(rrt, rt, 1920, 1080)
    assert np.array_equal(fb, fb2)                                  # idempotent
    assert (fb[0] == 0).all() and (hit[0] == -1).all() and (fb >> 24 == 0).all()
    monkeypatch.setenv("RRT_CULL_THRESHOLD", "0")
    rt0 = rrt.RayTracer(sd)
    # culling off = the same hierarchy code with infinite boxes: every record of a visited list is tested
    def mt_tests(r):
        c = rrt.abi.Counters()
        p = r.params(480, 270)
        rrt.abi.check(rrt.abi.lib().rrt_render_counters(r.handle, C.byref(p), None, C.byref(c)))
        return c.mt_tests
    assert mt_tests(rt0) > 20 * mt_tests(rt)
    fb0, hit0 = render(rrt, rt0, 1920, 1080, hits=True)
    assert np.array_equal(hit, hit0) and np.array_equal(fb, fb0)
    rt.close(); rt0.close()


def test_config4_sharded_equals_whole(rrt):
    # config[3] geometry (1M-triangle sphere grid) at 3840x2160: rendering interleaved tile rows on
    # 1/2/4/8 "ranks" (here sequentially on one GPU) and de-interleaving must give the whole frame.
    import torch
    from rrt_b200.synthetic import sphere_grid
    from rrt_b200.sharding import ShardedRenderer
    sd = sphere_grid()
    assert sd.n_tris == 999_936
    rt = rrt.RayTracer(sd)
    W, H = 3840, 2160
    whole, _ = render(rrt, rt, W, H)
    assert (whole[0] == 0).all() and (whole[1:] != 0x00FFFFFF).mean() > 0.2
    p = rt.params(W, H)
    L = rrt.abi.lib()
    for G in (2, 8):
        rs = [ShardedRenderer(rt, W, H, r, G, tile_rows=8) for r in range(G)]
        gathered = torch.zeros((G, rs[0].rpb, W), dtype=torch.int32, device="cuda:0")
        for r in rs:
            r.render(p)
            gathered[r.rank].copy_(r.local)
        frame = torch.zeros((H, W), dtype=torch.int32, device="cuda:0")
        rrt.abi.check(L.rrt_deinterleave_device(0, gathered.data_ptr(), rs[0].rpb, W, H, 8, G, frame.data_ptr(), None))
        torch.cuda.synchronize()
        assert np.array_equal(frame.cpu().numpy().view(np.uint32), whole)
    rt.close()


def test_counters_match_oracle_ray_counts(rrt, scenes):
    # the instrumented kernel counts the same rays the oracle traces (primary / shadow / reflection)
    from oracle.oracle import OracleScene
    sd = scenes("model2")
    rt = rrt.RayTracer(sd)
    p = rt.params(160, 90)
    c = rrt.abi.Counters()
    rrt.abi.check(rrt.abi.lib().rrt_render_counters(rt.handle, C.byref(p), None, C.byref(c)))
    _, _, _, oc = OracleScene(sd).render(p, counters=True)
    assert (c.rays_primary, c.rays_shadow, c.rays_reflect) == (oc["rays_primary"], oc["rays_shadow"], oc["rays_reflect"])
    assert 0 < c.mt_tests < oc["n_a"]          # culling: fewer exact tests than the reference executes
    rt.close()


def test_fma_quotient_equals_ieee_division(rrt):
    # the child-slab quotients use RN(1/d) + two FMA corrections instead of nine divisions per node;
    # every such quotient must be the IEEE quotient, bit for bit (2^31 operand pairs, hard mantissas included)
    n = C.c_uint64(0)
    rrt.abi.check(rrt.abi.lib().rrt_selftest_division(0, 1 << 31, C.byref(n)))
    assert n.value == 0


def test_fma_quotient_path_equals_division_path(rrt, scenes, monkeypatch):
    # same frame with the FMA quotient path disabled (true divisions everywhere): identical output
    sd = scenes("model2")
    rt = rrt.RayTracer(sd)
    fb, hit = render(rrt, rt, 960, 540, hits=True)
    monkeypatch.setenv("RRT_SLOW_DIV", "1")
    rt2 = rrt.RayTracer(sd)
    fb2, hit2 = render(rrt, rt2, 960, 540, hits=True)
    assert np.array_equal(fb, fb2) and np.array_equal(hit, hit2)
    rt.close(); rt2.close()


def test_config5_full_size_lattice_against_oracle(rrt):
    # config[4] at FULL size (10M-triangle soup, 7680x4320), orbit frame 5: the oracle cannot render the frame
    # (a ray scans an 82,834-triangle root list), so it renders a 96-pixel lattice of it; hit ids and pixels
    # must be identical there, and the frame must be idempotent.
    from oracle.oracle import OracleScene
    from rrt_b200.synthetic import triangle_soup
    sd = triangle_soup()
    rt = rrt.RayTracer(sd)
    st = rt.stats()
    assert st["n_tris_stored"] == 10_000_000 and st["root_list"] > 50_000
    W, H = 7680, 4320
    origin, rot = rrt.orbit_camera(5)
    sc = rrt.Scene(W, H)
    fb = sc.draw_scene(rt, want_hit_ids=True, origin=origin, camera_rot=rot).copy()
    hit = sc.hit_ids
    fb2 = sc.draw_scene(rt, origin=origin, camera_rot=rot)
    assert np.array_equal(fb, fb2) and (fb[0] == 0).all()
    step = 96
    ofb, ohit, _, _ = OracleScene(sd).render(rt.params(W, H, origin=origin, camera_rot=rot), row_step=step, col_step=step)
    rows, cols = np.arange(1, H, step), np.arange(0, W, step)
    assert np.array_equal(hit[np.ix_(rows, cols)], ohit[np.ix_(rows, cols)])
    assert np.array_equal(fb[np.ix_(rows, cols)], ofb[np.ix_(rows, cols)])
    assert (ohit[np.ix_(rows, cols)] >= 0).mean() > 0.9
    rt.close()


def test_fused_render_into_full_frame_equals_whole(rrt, scenes):
    # rrt_render_device_into_frame: every shard stores its rows straight into ONE full frame (what the ranks
    # do over NVLink peer memory at N > 1); here the shards run one after another on one GPU.
    import torch
    sd = scenes("model2")
    rt = rrt.RayTracer(sd)
    W, H = 1000, 562
    whole, _ = render(rrt, rt, W, H)
    p = rt.params(W, H)
    L = rrt.abi.lib()
    for G, T in ((2, 8), (3, 5), (8, 8)):
        fp = C.c_void_p()
        rrt.abi.check(L.rrt_frame_alloc(0, W, H, C.byref(fp)))
        for r in range(G):
            sh = rrt.abi.Shard(T, G, r)
            rrt.abi.check(L.rrt_render_device_into_frame(rt.handle, C.byref(p), C.byref(sh), fp, None))
        out = np.empty((H, W), dtype=np.uint32)
        rrt.abi.check(L.rrt_copy_frame_to_host(0, fp, W, H, out.ctypes.data, None))
        torch.cuda.synchronize()
        assert np.array_equal(out, whole)
        rrt.abi.check(L.rrt_frame_free(0, fp))
    rt.close()


def _peer_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    import rrt_b200
    from rrt_b200.loader import load_scene_npz
    from rrt_b200.sharding import ShardedRenderer
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    sd = load_scene_npz(os.path.join(os.path.dirname(__file__), "golden", "scenes", "model2.npz"))
    rt = rrt_b200.RayTracer(sd, device=rank)
    W, H = 640, 360
    # a MOVING camera: every frame differs, so a frame torn between two renders (a rank storing frame k+1 while rank 0
    # still reads frame k) cannot go unnoticed
    cams = []
    for k in (0, 3, 7, 12, 18, 25):
        origin, rot = rrt_b200.orbit_camera(k)
        cams.append(rt.params(W, H, origin=origin, camera_rot=rot))
    out = {}
    for mode in ("peer", "nccl", "host"):
        sr = ShardedRenderer(rt, W, H, rank, world, tile_rows=8, assemble=mode)
        frames = []
        host = torch.empty((len(cams), H, W), dtype=torch.int32, pin_memory=True)
        for i, p in enumerate(cams):
            sr.render(p)
            f = sr.assemble()
            if rank == 0:
                if mode == "host":
                    frames.append(np.array(f, dtype=np.uint32, copy=True))      # read before the next render() call
                else:
                    host[i].copy_(f, non_blocking=True)                           # consumed ON the render stream
        torch.cuda.synchronize()
        if rank == 0:
            out[mode] = frames if mode == "host" else [host[i].numpy().view(np.uint32).copy() for i in range(len(cams))]
        if rank != 0:
            sr.close()
        dist.barrier()
        if rank == 0:
            sr.close()
    if rank == 0:
        sc = rrt_b200.Scene(W, H)
        out["whole"] = [sc.draw_scene(rt, origin=tuple(p.origin), camera_rot=(np.array(p.camera_rot).reshape(3, 3) if p.use_camera_rot else None)).copy()
                        for p in cams]
        q.put(out)
    rt.close()
    dist.barrier()
    dist.destroy_process_group()


def test_peer_frame_two_gpus(rrt):
    # N = 2, one process per GPU, a moving camera: rank 1's render kernel stores into rank 0's frames through CUDA IPC /
    # NVLink (two frames used alternately), the NCCL all-gather mode, and the shared page-locked HOST frame every rank
    # copies its rows into — each of the six frames must equal the single-GPU frame of the same camera
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_peer_worker, args=(r, 2, 29541, q)) for r in range(2)]
    for pr in procs:
        pr.start()
    out = q.get(timeout=300)
    for pr in procs:
        pr.join(timeout=120)
        assert pr.exitcode == 0
    assert any(not np.array_equal(out["whole"][0], w) for w in out["whole"][1:])
    for mode in ("peer", "nccl", "host"):
        for i, w in enumerate(out["whole"]):
            assert np.array_equal(out[mode][i], w), (mode, i)


def test_host_frame_single_rank_equals_rrt_render(rrt, scenes):
    # rrt_render_to_host_frame with one shard, and with the shards of a 3-way split rendered one after another into one
    # host frame (pitched copies of whole tile-row groups + the ragged last group): the frame rrt_render gives
    import torch
    sd = scenes("model2")
    rt = rrt.RayTracer(sd)
    L = rrt.abi.lib()
    for W, H, G, T in ((640, 360, 1, 8), (333, 101, 3, 8), (200, 57, 2, 5)):
        p = rt.params(W, H)
        whole = rrt.Scene(W, H).draw_scene(rt).copy()
        host = torch.zeros((H, W), dtype=torch.int32, pin_memory=True)
        for r in range(G):
            sh = rrt.abi.Shard(T, G, r)
            rrt.abi.check(L.rrt_render_to_host_frame(rt.handle, C.byref(p), C.byref(sh), host.data_ptr(), None))
            torch.cuda.synchronize()
        assert np.array_equal(host.numpy().view(np.uint32), whole), (W, H, G, T)
    rt.close()


def test_page_locked_host_buffers_take_direct_stores_and_give_the_same_frame(rrt, scenes):
    # A page-locked caller buffer is written by the render kernel itself (posted PCIe stores, no device->host copy:
    # host_frame_alias in csrc/rrt_api.cu); a pageable one goes through render + copy.  Same frame either way, for
    # rrt_render and for the shards of rrt_render_to_host_frame, including buffers at an offset inside a pinned block.
    import torch
    sd = scenes("model2")
    rt = rrt.RayTracer(sd)
    L = rrt.abi.lib()
    for W, H, G, T in ((320, 180, 1, 8), (333, 101, 3, 8)):
        p = rt.params(W, H)
        pageable = np.full((H, W), 0xDEADBEEF, dtype=np.uint32)
        rrt.abi.check(L.rrt_render(rt.handle, C.byref(p), pageable.ctypes.data_as(C.c_void_p), None, None))
        block = torch.full((H * W + 64,), -1, dtype=torch.int32, pin_memory=True)
        pinned = block[37:37 + H * W]                       # not at the start of the allocation
        rrt.abi.check(L.rrt_render(rt.handle, C.byref(p), C.c_void_p(pinned.data_ptr()), None, None))
        assert np.array_equal(pinned.numpy().view(np.uint32).reshape(H, W), pageable), (W, H)
        assert (block[:37] == -1).all() and (block[37 + H * W:] == -1).all()    # nothing outside the frame was touched
        for host in (torch.zeros((H, W), dtype=torch.int32, pin_memory=True), np.zeros((H, W), dtype=np.uint32)):
            ptr = host.data_ptr() if hasattr(host, "data_ptr") else host.ctypes.data
            for r in range(G):
                sh = rrt.abi.Shard(T, G, r)
                rrt.abi.check(L.rrt_render_to_host_frame(rt.handle, C.byref(p), C.byref(sh), C.c_void_p(ptr), None))
                torch.cuda.synchronize()
            got = host.numpy().view(np.uint32) if hasattr(host, "data_ptr") else host
            assert np.array_equal(got, pageable), (W, H, G, T, type(host))
    rt.close()


def test_reserve_changes_nothing_but_where_the_allocation_happens(rrt, scenes):
    # rrt_scene_reserve: the first frame of a reserved shape, of a smaller and of a LARGER shape than reserved, whole
    # and sharded — always the frame an unreserved handle renders
    sd = scenes("model2")
    plain, res = rrt.RayTracer(sd), rrt.RayTracer(sd)
    L = rrt.abi.lib()
    p0 = res.params(640, 360)
    rrt.abi.check(L.rrt_scene_reserve(res.handle, C.byref(p0), None, 1))
    sh = rrt.abi.Shard(8, 3, 1)
    rrt.abi.check(L.rrt_scene_reserve(res.handle, C.byref(p0), C.byref(sh), 0))
    for W, H in ((640, 360), (320, 200), (1280, 720), (640, 360)):
        a = rrt.Scene(W, H).draw_scene(plain).copy()
        sc = rrt.Scene(W, H)
        if (W, H) == (1280, 720):
            sc.reserve(res)
        assert np.array_equal(sc.draw_scene(res), a), (W, H)
    bad = res.params(640, 360)
    bad.max_depth = 99
    assert L.rrt_scene_reserve(res.handle, C.byref(bad), None, 0) != 0      # same argument checks as a render call
    plain.close()
    res.close()


def test_order_hint_changes_nothing(rrt, scenes):
    # rrt_scene_order_hint: from the second frame of a shape on, the last part of the frame's 16x4-pixel blocks is
    # handed out heaviest first (cost measured during the previous frame, sorted on a side stream).  Only the ORDER
    # of the work may change: frames and hit ids must be bit-identical with the hint off, on, after a change of
    # shape, and when shards of different shapes alternate on one handle.
    import torch
    sd = scenes("model2")
    rt = rrt.RayTracer(sd)
    L = rrt.abi.lib()
    rrt.abi.check(L.rrt_scene_order_hint(rt.handle, 0))
    ref = {}
    for W, H in ((1920, 1080), (1000, 562)):
        ref[(W, H)] = render(rrt, rt, W, H, hits=True)
    rrt.abi.check(L.rrt_scene_order_hint(rt.handle, 1))
    for W, H in ((1920, 1080), (1920, 1080), (1920, 1080), (1000, 562), (1000, 562), (1920, 1080), (1920, 1080)):
        fb, hit = render(rrt, rt, W, H, hits=True)
        assert np.array_equal(fb, ref[(W, H)][0]) and np.array_equal(hit, ref[(W, H)][1])
    # shards of one frame, alternating on the same handle, each rendered three times into one full frame
    W, H = 1920, 1080
    p = rt.params(W, H)
    fp = C.c_void_p()
    rrt.abi.check(L.rrt_frame_alloc(0, W, H, C.byref(fp)))
    for rep in range(3):
        for r in range(2):
            sh = rrt.abi.Shard(8, 2, r)
            for _ in range(3 if rep == 1 else 1):
                rrt.abi.check(L.rrt_render_device_into_frame(rt.handle, C.byref(p), C.byref(sh), fp, None))
        out = np.empty((H, W), dtype=np.uint32)
        rrt.abi.check(L.rrt_copy_frame_to_host(0, fp, W, H, out.ctypes.data, None))
        torch.cuda.synchronize()
        assert np.array_equal(out, ref[(W, H)][0])
    rrt.abi.check(L.rrt_frame_free(0, fp))
    rt.close()


def test_render_frames_equals_single_frames(rrt, scenes):
    # rrt_render_frames: frame i is copied to the host while frame i+1 renders (two device frames, render + copy
    # stream).  Every frame of the sequence must be what rrt_render gives for the same params — with pageable and
    # with pinned host memory, and with more frames than device buffers.
    import torch
    sd = scenes("model2")
    rt = rrt.RayTracer(sd)
    W, H = 640, 360
    cams = []
    for k in (0, 5, 11, 16, 23):
        origin, rot = rrt.orbit_camera(k)
        cams.append(dict(origin=origin, camera_rot=rot))
    sc = rrt.Scene(W, H)
    singles = [sc.draw_scene(rt, **c).copy() for c in cams]
    assert any(not np.array_equal(singles[0], s) for s in singles[1:])
    seq = rrt.Scene(W, H).draw_frames(rt, cams)
    pinned = torch.empty((len(cams), H, W), dtype=torch.int32, pin_memory=True)
    seq_pinned = rrt.Scene(W, H).draw_frames(rt, cams, out=pinned.numpy().view(np.uint32))
    for i, s in enumerate(singles):
        assert np.array_equal(seq[i], s) and np.array_equal(seq_pinned[i], s)
    assert rrt.Scene(W, H).draw_frames(rt, []).shape == (0, H, W)
    rt.close()


def test_frame_parallel_sequence_world1(rrt, scenes):
    # frame-parallel sequences (sharding.FrameParallelSequence): with one rank the sequence is what draw_frames gives;
    # the N>1 assignment / gather is covered on CPU (tests/test_sharding_gloo.py)
    from rrt_b200.sharding import FrameParallelSequence
    sd = scenes("model2")
    rt = rrt.RayTracer(sd)
    W, H = 320, 180
    cams = []
    for k in (0, 9, 31):
        origin, rot = rrt.orbit_camera(k)
        cams.append(dict(origin=origin, camera_rot=rot))
    fps = FrameParallelSequence(rt, W, H, rank=0, world=1)
    local = fps.render_local(cams)
    seq = fps.collect(local, len(cams))
    sc = rrt.Scene(W, H)
    for i, c in enumerate(cams):
        assert np.array_equal(seq[i], sc.draw_scene(rt, **c))
    rt.close()


def test_max_depth_beyond_the_mirror_stack_is_refused(rrt, scenes):
    # the reference recurses to any depth (raytracer.rs:76); the kernel keeps 16 mirror levels per lane: a deeper
    # request must fail loudly (RRT_ERR_INVALID), never render a different image with RRT_OK
    rt = rrt.RayTracer(scenes("model2"))
    sc = rrt.Scene(32, 16)
    sc.draw_scene(rt, max_depth=16)
    with pytest.raises(Exception) as ei:
        sc.draw_scene(rt, max_depth=17)
    assert "max_depth" in str(ei.value)
    rt.close()
