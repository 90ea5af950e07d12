"""Regenerates tests/golden/frames/*.npz: small frames (packed 0x00RRGGBB buffer + depth-0 hit ids) and
seeded per-ray results produced by the CPU oracle (oracle/rrt_oracle.cpp) on the scene fixtures.
The reference itself cannot run here (no Rust toolchain), so these vectors pin the ORACLE's behaviour
over time and give the GPU tests a fixture that does not need liboracle at all."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import rrt_b200  # noqa: E402
from rrt_b200.loader import load_scene_npz  # noqa: E402
from oracle.oracle import OracleScene  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "frames")
W, H = 128, 72


def golden_rays(seed, n=4096):
    rng = np.random.default_rng(seed)
    org = np.tile(np.array([0.0, 2.0, -10.0]), (n, 1))
    org[n // 2:] = rng.uniform(-6, 6, (n - n // 2, 3)) + np.array([0.0, 2.0, 0.0])
    d = np.concatenate([rng.uniform(-0.5, 0.5, (n, 2)), np.ones((n, 1))], 1)
    d[n // 2:] = rng.normal(size=(n - n // 2, 3))
    d[::13, 0] = 0.0
    mt = np.where(rng.random(n) < 0.5, np.inf, rng.uniform(1.0, 30.0, n))
    return org, d, mt


def main():
    os.makedirs(OUT, exist_ok=True)
    for name in ("model", "model2", "model3"):
        sd = load_scene_npz(os.path.join(ROOT, "tests", "golden", "scenes", name + ".npz"))
        o = OracleScene(sd)
        fb, hit, _, cnt = o.render(rrt_b200.make_params(W, H), counters=True)
        org, d, mt = golden_rays(seed=len(name))
        tri, tuv = o.trace(org, d, mt)
        rgb = o.ray_colours(rrt_b200.make_params(W, H), org[:1024], d[:1024])
        np.savez_compressed(os.path.join(OUT, name + ".npz"), fb=fb, hit=hit, tri=tri, tuv=tuv, rgb=rgb,
                            counters=np.array([cnt[k] for k in sorted(cnt)], dtype=np.uint64))
        print(name, "hits", int((hit[1:] >= 0).sum()), "of", hit[1:].size, sorted(cnt))


if __name__ == "__main__":
    main()
