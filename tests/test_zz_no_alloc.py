"""rrt_render_device_noalloc — the allocation-free variant of rrt_render_device
(no counterpart in the reference): its scratch is a static device array inside the library instead of a
cudaMalloc'ed buffer.  Checked through the C ABI against the existing entry point, bit for bit (packed integer
pixels and hit ids), on calls in a row that differ from one another — the array keeps the previous call's contents,
so a call that failed to clear what it uses would show up in the next frame.

Named test_zz_* so that pytest collects it LAST: a failure here must not hide
any other test.  Passes alone in a fresh process and after the other tests (it builds its own scene handle)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _render(fn, rt, p, shard, rows, W, want_hits):
    """One call of `fn` (rrt_render_device or rrt_render_device_noalloc) into fresh device buffers -> host arrays."""
    import torch
    from rrt_b200 import abi
    dev = torch.device("cuda", rt.device)
    fb = torch.full((rows, W), 0x5A5A5A5A, dtype=torch.int32, device=dev)        # garbage: every pixel must be written
    hit = torch.full((rows, W, 4), -7, dtype=torch.int32, device=dev) if want_hits else None
    abi.check(fn(rt.handle, C.byref(p), None if shard is None else C.byref(shard), fb.data_ptr(),
                 None if hit is None else hit.data_ptr(), None))
    torch.cuda.synchronize(dev)
    return fb.cpu().numpy().view(np.uint32), (None if hit is None else hit.cpu().numpy())


def test_noalloc_entry_point_equals_rrt_render_device(rrt, scenes):
    L = rrt.abi.lib()
    sd = scenes("model2")                      # bump maps, shadows, mirror bounces: every kind of ray
    rt = rrt.RayTracer(sd)
    # (size, camera) of consecutive calls: a large frame, then smaller and different ones, then the first again
    calls = [(640, 360, 0), (320, 200, 9), (333, 101, 17), (640, 360, 0)]
    for W, H, k in calls:
        origin, rot = rrt.orbit_camera(k)
        p = rt.params(W, H, origin=origin, camera_rot=rot)
        ref_fb, ref_hit = _render(L.rrt_render_device, rt, p, None, H, W, True)
        fb, hit = _render(L.rrt_render_device_noalloc, rt, p, None, H, W, True)
        assert np.array_equal(fb, ref_fb), (W, H, k)
        assert np.array_equal(hit, ref_hit), (W, H, k)
        assert (fb[0] == 0).all() and (fb[1:] != 0x5A5A5A5A).all()
        fb2, _ = _render(L.rrt_render_device_noalloc, rt, p, None, H, W, False)       # same call again, no hit ids
        assert np.array_equal(fb2, ref_fb), (W, H, k)
    # a shard (interleaved tile rows, compact output) through both entry points
    W, H = 640, 360
    p = rt.params(W, H)
    for r in range(3):
        sh = rrt.abi.Shard(8, 3, r)
        rows = int(L.rrt_shard_rows(H, C.byref(sh)))
        ref_fb, _ = _render(L.rrt_render_device, rt, p, sh, rows, W, False)
        fb, _ = _render(L.rrt_render_device_noalloc, rt, p, sh, rows, W, False)
        assert np.array_equal(fb, ref_fb), r
    rt.close()


def test_noalloc_entry_point_refuses_what_does_not_fit(rrt, scenes):
    import torch
    L = rrt.abi.lib()
    rt = rrt.RayTracer(scenes("model2"))
    assert 1920 * 1080 <= rrt.abi.RRT_NOALLOC_MAX_PIXELS < 3840 * 2160
    p = rt.params(3840, 2160)
    dummy = torch.zeros(16, dtype=torch.int32, device="cuda")     # never written: the call is refused before any launch
    rc = L.rrt_render_device_noalloc(rt.handle, C.byref(p), None, dummy.data_ptr(), None, None)
    assert rc == -1                                               # RRT_ERR_INVALID
    msg = L.rrt_last_error().decode()
    assert "RRT_NOALLOC_MAX_PIXELS" in msg and str(3840 * 2160) in msg and str(rrt.abi.RRT_NOALLOC_MAX_PIXELS) in msg
    # the same frame in 8 shards fits (1 036 800 pixels each) — and the handle still renders after the refusal
    sh = rrt.abi.Shard(8, 8, 3)
    rows = int(L.rrt_shard_rows(2160, C.byref(sh)))
    assert rows * 3840 <= rrt.abi.RRT_NOALLOC_MAX_PIXELS
    ref_fb, _ = _render(L.rrt_render_device, rt, p, sh, rows, 3840, False)
    fb, _ = _render(L.rrt_render_device_noalloc, rt, p, sh, rows, 3840, False)
    assert np.array_equal(fb, ref_fb)
    rt.close()
