"""A host with no Python and no torch.distributed drives N ranks through the C ABI alone: tests/native/multi_rank_driver.cpp
is compiled against include/rrt.h, linked with librrt_b200.so only, and launched as `world` processes that assemble
frames of a moving camera in ONE shared page-locked host canvas (rrt_host_frame_*, rrt_render_device,
rrt_copy_shard_to_host_frame, rrt_synchronize).  Rank 0 compares every frame with rrt_render's, word for word."""
import os
import subprocess

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

SRC = os.path.join(ROOT, "tests", "native", "multi_rank_driver.cpp")
EXE = os.path.join(ROOT, "tests", "native", "multi_rank_driver")
LIBDIR = os.path.join(ROOT, "rust-ray-tracer_b200", "csrc")


def _build():
    lib = os.path.join(LIBDIR, "librrt_b200.so")
    if os.path.exists(EXE) and os.path.getmtime(EXE) >= max(os.path.getmtime(SRC), os.path.getmtime(lib)):
        return
    subprocess.run(["g++", "-O2", "-std=c++17", "-Wall", "-o", EXE, SRC, "-L" + LIBDIR, "-lrrt_b200",
                    "-Wl,-rpath," + LIBDIR], check=True)


@pytest.mark.parametrize("world,W,H", [(2, 320, 180), (3, 200, 117)])
def test_native_ranks_assemble_frames_in_a_shared_host_canvas(rrt, world, W, H):
    _build()
    n_dev = rrt.abi.lib().rrt_device_count()
    assert n_dev >= 1
    name = f"/rrt_native_{os.getpid()}_{world}"
    procs = [subprocess.Popen([EXE, str(r), str(world), name, str(W), str(H), str(r % n_dev), "4"],
                              stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True) for r in range(world)]
    outs = []
    try:
        for p in procs:
            outs.append(p.communicate(timeout=240))
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    for r, (p, (out, err)) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r} rc {p.returncode}\n{out}\n{err}"
    assert "ALL FRAMES EQUAL" in outs[0][0], outs[0][0]
    assert outs[0][0].count(" 0 of ") == 4, outs[0][0]
    assert not os.path.exists("/dev/shm" + name)
