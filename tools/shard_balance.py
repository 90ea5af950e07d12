"""Development aid: time every shard of a G-way row-interleaved frame on ONE GPU for several tile_rows,
to pick the interleave that balances the ranks (max over shards is what an N-GPU frame costs)."""
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
Ts = sys.argv[3:] or ["1", "2", "4", "8", "16"]
desc, W, H = bench.WORKLOADS[name]
rt = rrt_b200.RayTracer(bench.load_workload(name))
p = rt.params(W, H)
L = abi.lib()
fb = torch.zeros((H, W), dtype=torch.int32, device="cuda:0")
flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda:0")
for T in Ts:
    ms = []
    for r in range(G):
        sh = abi.Shard(int(T), G, r)
        best = 1e9
        for it in range(4):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            abi.check(L.rrt_render_device_into_frame(rt.handle, C.byref(p), C.byref(sh), fb.data_ptr(), None))
            e1.record()
            torch.cuda.synchronize()
            if it:
                best = min(best, e0.elapsed_time(e1))
        ms.append(best)
    print(f"{name} G={G} tile_rows={T}: max {max(ms):.3f} mean {sum(ms) / G:.3f} min {min(ms):.3f} ms  "
          f"balance {sum(ms) / G / max(ms):.3f}  per shard {[round(x, 2) for x in ms]}", flush=True)
