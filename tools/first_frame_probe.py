"""Development aid: what the FIRST frame of a shape costs on every shard of a G-way split (one GPU, shards one after
another), next to the steady state with and without the dispatch-order hint.  (RRT_ORDER_ESTIMATE selected the probe-ray
cost estimate of profiles/r02_experiments.md section 6 while that experiment was in the tree; it is ignored now.)
  python tools/first_frame_probe.py [c4] [8]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
import rrt_b200  # noqa: E402
from rrt_b200 import abi  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c4"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 8
desc, W, H = bench.WORKLOADS[name]
sd = bench.load_workload(name)
L = abi.lib()
fb = torch.zeros((H, W), dtype=torch.int32, device="cuda:0")
flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda:0")


def timed(rt, p, sh):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    abi.check(L.rrt_render_device_into_frame(rt.handle, C.byref(p), C.byref(sh), fb.data_ptr(), None))
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


ref = None
for est in ("1", "0"):
    os.environ["RRT_ORDER_ESTIMATE"] = est
    rt = rrt_b200.RayTracer(sd)
    p = rt.params(W, H)
    first, steady, nohint = [], [], []
    for r in range(G):
        sh = abi.Shard(8, G, r)
        abi.check(L.rrt_scene_reserve(rt.handle, C.byref(p), C.byref(sh), 0))
        timed(rt, p, abi.Shard(8, G, (r + 1) % G))          # another shape in between: this shard starts fresh
        f = []
        for _ in range(3):
            abi.check(L.rrt_scene_order_hint(rt.handle, 1))  # forget the schedule: the next frame is a first frame
            f.append(timed(rt, p, sh))
        first.append(min(f))
        for _ in range(2):
            timed(rt, p, sh)
        steady.append(min(timed(rt, p, sh) for _ in range(3)))
        abi.check(L.rrt_scene_order_hint(rt.handle, 0))
        timed(rt, p, sh)
        nohint.append(min(timed(rt, p, sh) for _ in range(3)))
        abi.check(L.rrt_scene_order_hint(rt.handle, 1))
    torch.cuda.synchronize()
    digest = int(fb.to(torch.int64).sum().item())
    if ref is None:
        ref = digest
    print(f"{name} G={G} estimate={est}: first frame max {max(first):.3f} mean {sum(first) / G:.3f} | hinted steady max {max(steady):.3f} "
          f"| no hint max {max(nohint):.3f} ms | frame sum {'same' if digest == ref else 'DIFFERENT'}", flush=True)
    print("   first per shard", [round(x, 3) for x in first], flush=True)
    rt.close()
