"""rust-ray-tracer_b200 — B200-native (sm_100a) drop-in for the per-pixel hot path of
conor722/rust-ray-tracer.  See DESIGN.md; C ABI in include/rrt.h; CUDA in csrc/."""
from . import abi  # noqa: F401
from .scene import (Canvas, Light, MaterialSpec, RayTracer, Scene, SceneData,  # noqa: F401
                    default_lights, make_params, orbit_camera, DEFAULT_ORIGIN)
