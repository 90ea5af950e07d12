"""Deterministic synthetic scenes of BASELINE.json configs 4 and 5. Vertices are generated
in f64 and handed to both the CUDA path and the oracle as binary arrays."""
import numpy as np

from .scene import MaterialSpec, SceneData


def checker_texture(size=256):
    """Lossless stand-in for map_test.png (256x256 RGB) so the GPU box needs no reference asset."""
    y, x = np.mgrid[0:size, 0:size]
    c = ((x // 32 + y // 32) & 1).astype(np.uint8)
    t = np.zeros((size, size, 3), dtype=np.uint8)
    t[..., 0] = 60 + 180 * c
    t[..., 1] = (x * 255 // (size - 1)).astype(np.uint8)
    t[..., 2] = (y * 255 // (size - 1)).astype(np.uint8)
    return t


def _sphere(cx, cy, cz, radius, nlon, nlat):
    """Lat-long sphere: 2*nlon*(nlat-1) triangles, per-vertex unit radial normals, vt=(lon/2pi*4, lat/pi*4)."""
    lat = np.linspace(0.0, np.pi, nlat + 1)
    lon = np.linspace(0.0, 2.0 * np.pi, nlon + 1)
    LA, LO = np.meshgrid(lat, lon, indexing="ij")          # (nlat+1, nlon+1)
    n = np.stack([np.sin(LA) * np.cos(LO), np.cos(LA), np.sin(LA) * np.sin(LO)], -1)
    p = n * radius + np.array([cx, cy, cz])
    t = np.stack([LO / (2 * np.pi) * 4.0, LA / np.pi * 4.0], -1)
    v, vt, vn = [], [], []

    def tri(a, b, c):
        v.append(np.concatenate([p[a], p[b], p[c]], -1))
        vt.append(np.concatenate([t[a], t[b], t[c]], -1))
        vn.append(np.concatenate([n[a], n[b], n[c]], -1))
    i, j = np.mgrid[0:nlat, 0:nlon]
    a, b, c, d = (i, j), (i + 1, j), (i + 1, j + 1), (i, j + 1)
    up = i > 0           # skip the degenerate triangle at the north pole
    dn = i < nlat - 1    # ... and at the south pole

    def sel(idx, m):
        return (idx[0][m], idx[1][m])
    tri(sel(a, dn), sel(b, dn), sel(c, dn))
    tri(sel(a, up), sel(c, up), sel(d, up))
    return np.concatenate(v), np.concatenate(vt), np.concatenate(vn)


def sphere_grid(nx=8, ny=8, nlon=126, nlat=63, radius=0.9, kr=0.0, texture=None):
    """Config 4: nx x ny spheres, centres x in {-7,-5,..,7}, y in {2-7,..,2+7}, z=8.
    Defaults give 64 * 2*126*62 = 999 936 triangles."""
    vs, vts, vns = [], [], []
    for iy in range(ny):
        for ix in range(nx):
            cx = -(nx - 1) + 2.0 * ix
            cy = 2.0 - (ny - 1) + 2.0 * iy
            v, vt, vn = _sphere(cx, cy, 8.0, radius, nlon, nlat)
            vs.append(v); vts.append(vt); vns.append(vn)
    v, vt, vn = np.concatenate(vs), np.concatenate(vts), np.concatenate(vns)
    tex = checker_texture() if texture is None else texture
    mat = MaterialSpec("sphere", (1.0, 1.0, 1.0), (1.0, 1.0, 1.0), (1.0, 1.0, 1.0), 240.0, kr, 0, -1)
    return SceneData(v, vt, vn, np.zeros(len(v), np.int32), [mat], [tex], texture_names=["checker"])


def triangle_soup(n=10_000_000, seed=12345, extent=18.0, size=0.1, kr=0.0, texture=None):
    """Config 5: n triangles, centre ~ U[-extent,extent]^3, vertices = centre + U[-size,size]^3,
    geometric normals, vt ~ U[0,1]^2."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-extent, extent, size=(n, 1, 3))
    p = c + rng.uniform(-size, size, size=(n, 3, 3))
    vt = rng.uniform(0.0, 1.0, size=(n, 6))
    e1, e2 = p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]
    g = np.cross(e1, e2)
    ln = np.linalg.norm(g, axis=1, keepdims=True)
    g = g / np.where(ln > 0, ln, 1.0)
    vn = np.concatenate([g, g, g], axis=1)
    tex = checker_texture() if texture is None else texture
    mat = MaterialSpec("soup", (1.0, 1.0, 1.0), (1.0, 1.0, 1.0), (1.0, 1.0, 1.0), 240.0, kr, 0, -1)
    return SceneData(p.reshape(n, 9), vt, vn, np.zeros(n, np.int32), [mat], [tex], texture_names=["checker"])
