"""Ad-hoc GPU parity / timing probe (development aid; the real gates are tests/ -m gpu and bench.py)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import rrt_b200  # noqa: E402
from rrt_b200 import RayTracer, Scene, make_params  # noqa: E402
from rrt_b200.loader import load_scene_npz  # noqa: E402
from rrt_b200.synthetic import sphere_grid, triangle_soup  # noqa: E402
from oracle.oracle import OracleScene  # noqa: E402


def compare(name, sd, W, H, oracle=True, **kw):
    rt = RayTracer(sd)
    print(name, "stats", rt.stats(), flush=True)
    sc = Scene(W, H)
    fb = sc.draw_scene(rt, want_hit_ids=True, **kw).copy()
    hit = sc.hit_ids
    ms = []
    for _ in range(3):
        sc.draw_scene(rt, **kw)
        ms.append(sc.device_ms)
    print(f"{name} {W}x{H}: device ms {ms} -> {4 * W * H / min(ms) / 1e3:.1f} Mrays/s", flush=True)
    if oracle:
        p = rt.params(W, H, **kw)
        osc = OracleScene(sd)
        ofb, ohit, sec, _ = osc.render(p)
        nh = int((hit != ohit).sum())
        d = np.abs(((fb[..., None] >> np.array([16, 8, 0])) & 255).astype(int) - ((ofb[..., None] >> np.array([16, 8, 0])) & 255).astype(int))
        print(f"{name}: oracle {sec:.2f}s ({4*W*(H-1)/sec/1e6:.3f} Mrays/s) hit mismatches {nh}/{hit.size} "
              f"fb exact-diff pixels {int((fb != ofb).sum())} max channel diff {int(d.max())} row0 zero {bool((fb[0] == 0).all())}",
              flush=True)
        if nh:
            idx = np.argwhere(hit != ohit)[:10]
            for r, c, k in idx:
                print("   mismatch at", r, c, k, "gpu", hit[r, c, k], "oracle", ohit[r, c, k])
    rt.close()


if __name__ == "__main__":
    gold = os.path.join(ROOT, "tests", "golden", "scenes")
    which = sys.argv[1:] or ["model2s", "model2", "model", "model3", "c4"]
    for w in which:
        if w == "model2s":
            compare("model2-small", load_scene_npz(os.path.join(gold, "model2.npz")), 320, 180)
        elif w in ("model", "model2", "model3"):
            compare(w, load_scene_npz(os.path.join(gold, w + ".npz")), 1920, 1080, oracle=os.environ.get("FULL_ORACLE", "0") == "1")
        elif w == "c4s":
            compare("c4-small", sphere_grid(nlon=32, nlat=16), 480, 270)
        elif w == "c5s":
            sd = triangle_soup(20000, size=0.4)
            for k in (0, 5, 23):
                origin, rot = rrt_b200.orbit_camera(k)
                compare(f"c5-small frame {k}", sd, 320, 180, origin=origin, camera_rot=rot)
        elif w == "c5":
            t0 = time.time()
            sd = triangle_soup()
            print("c5 gen", time.time() - t0, sd.n_tris, flush=True)
            t0 = time.time()
            rt = RayTracer(sd)
            print("c5 scene create", time.time() - t0, rt.stats(), flush=True)
            W, H = 7680, 4320
            sc = Scene(W, H)
            osc = OracleScene(sd)
            for k in (0, 5):
                origin, rot = rrt_b200.orbit_camera(k)
                fb = sc.draw_scene(rt, want_hit_ids=True, origin=origin, camera_rot=rot).copy()
                print(f"c5 frame {k}: device ms {sc.device_ms:.2f} -> {4 * W * H / sc.device_ms / 1e3:.1f} Mrays/s", flush=True)
                p = rt.params(W, H, origin=origin, camera_rot=rot)
                st = 96
                ofb, ohit, sec, _ = osc.render(p, row_step=st, col_step=st)
                rows, cols = np.arange(1, H, st), np.arange(0, W, st)   # the oracle lattice starts at buffer row 1
                g = sc.hit_ids[np.ix_(rows, cols)]
                o = ohit[np.ix_(rows, cols)]
                gfb, ofbs = fb[np.ix_(rows, cols)], ofb[np.ix_(rows, cols)]
                print(f"   oracle lattice {len(rows)}x{len(cols)} in {sec:.1f}s: hit mismatches {(g != o).sum()} / {g.size}, "
                      f"hits {(o >= 0).sum()}, fb diff pixels {(gfb[1:] != ofbs[1:]).sum()}", flush=True)
            rt.close()
        elif w == "c4":
            t0 = time.time()
            sd = sphere_grid()
            print("c4 gen", time.time() - t0, sd.n_tris)
            compare("c4", sd, 3840, 2160, oracle=False)
