"""Development aid: time differently-built copies of librrt_b200.so on the GPU box and check that every
variant produces the SAME frame and hit-id plane (sha1) as the first one.

  build here :  RRT_OUT=$PWD/variants/v_name.so RRT_EXTRA_FLAGS="-DFOO=1" python rust-ray-tracer_b200/build.py --force
  on the box :  python tools/variant_sweep.py variants/*.so [-- c2 c4]
"""
import hashlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def one(workloads):
    sys.path.insert(0, ROOT)
    import bench
    import rrt_b200
    for name in workloads:
        desc, W, H = bench.WORKLOADS[name]
        rt = rrt_b200.RayTracer(bench.load_workload(name))
        sc = rrt_b200.Scene(W, H)
        fb = sc.draw_scene(rt, want_hit_ids=True).copy()
        h = hashlib.sha1(fb.tobytes()).hexdigest()[:12] + "/" + hashlib.sha1(sc.hit_ids.tobytes()).hexdigest()[:12]
        ms = []
        for _ in range(6):
            sc.draw_scene(rt)
            ms.append(sc.device_ms)
        print(f"   {name}: {min(ms):8.3f} ms  {4 * W * H / min(ms) / 1e3:8.1f} Mrays/s  sha {h}", flush=True)
        rt.close()


if __name__ == "__main__":
    args = sys.argv[1:]
    if args and args[0] == "--one":
        one(args[1:])
    else:
        wl = ["c2", "c4"]
        if "--" in args:
            wl = args[args.index("--") + 1:]
            args = args[:args.index("--")]
        for lib in args:
            print("==", lib, flush=True)
            env = dict(os.environ, RRT_LIB_PATH=os.path.abspath(lib))
            subprocess.run([sys.executable, os.path.abspath(__file__), "--one"] + wl, env=env)
