"""Second, independent restatement of the reference path in pure Python floats (IEEE f64, no FMA) —
TEST INFRASTRUCTURE, small cases only.  It exists to cross-check oracle/rrt_oracle.cpp: two
transcriptions written separately from the same Rust source must agree bit-for-bit.
Citations are to /root/reference/src/..."""
import math

EPS = 2.220446049250313e-16  # f64::EPSILON


def _fmin(a, b):  # Rust f64::min: NaN loses
    if a != a:
        return b
    if b != b:
        return a
    return a if a < b else b


def _fmax(a, b):
    if a != a:
        return b
    if b != b:
        return a
    return a if a > b else b


def _div(a, b):  # IEEE division incl. /0
    try:
        return a / b
    except ZeroDivisionError:
        if a != a or a == 0.0:
            return math.nan
        neg = (math.copysign(1.0, a) < 0) != (math.copysign(1.0, b) < 0)
        return -math.inf if neg else math.inf


def sub(a, b): return (a[0] - b[0], a[1] - b[1], a[2] - b[2])
def add(a, b): return (a[0] + b[0], a[1] + b[1], a[2] + b[2])
def mul(a, s): return (a[0] * s, a[1] * s, a[2] * s)
def divs(a, s): return (_div(a[0], s), _div(a[1], s), _div(a[2], s))
def dot(a, b): return (a[0] * b[0]) + (a[1] * b[1]) + (a[2] * b[2])              # engine.rs:85-87
def length(a): return math.sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2])         # engine.rs:89-91
def cross(a, b): return (a[1] * b[2] - a[2] * b[1], -(a[0] * b[2] - a[2] * b[0]), a[0] * b[1] - a[1] * b[0])
def normalised(a): return divs(a, length(a))


class Octree:
    """octree.rs:15-241. aabbs hold (min, max) tuples for node boxes and triangle boxes alike."""

    def __init__(self, mn, mx):
        self.aabbs = [(tuple(mn), tuple(mx))]
        self.nodes = [dict(aabb=0, tris=[], kids=[], count=0)]
        self.tris = []

    @staticmethod
    def overlap(a, b):  # aabb.rs:49-60
        for k in range(3):
            if a[1][k] < b[0][k] or a[0][k] > b[1][k]:
                return False
        return True

    def push(self, v1, v2, v3):  # octree.rs:41-52
        mn = tuple(_fmin(v1[k], _fmin(v2[k], v3[k])) for k in range(3))
        mx = tuple(_fmax(v1[k], _fmax(v2[k], v3[k])) for k in range(3))
        ai, ti = len(self.aabbs), len(self.tris)
        self.tris.append((v1, v2, v3))
        self.aabbs.append((mn, mx))
        self._push_at(ti, ai, 0)

    def _push_at(self, ti, ai, oi):  # octree.rs:54-108
        node = self.nodes[oi]
        if not self.overlap(self.aabbs[ai], self.aabbs[node["aabb"]]):
            return
        has, kids = bool(node["tris"]), list(node["kids"])
        leaf = not kids
        node["count"] += 1
        if leaf and not has:
            node["tris"].append(ti)
            return
        if leaf:
            kids = self._subdivide(oi)
        hit = [c for c in kids if self.overlap(self.aabbs[self.nodes[c]["aabb"]], self.aabbs[ai])]
        if len(hit) == 1:
            self._push_at(ti, ai, hit[0])
        else:
            node["tris"].append(ti)

    def _subdivide(self, oi):  # octree.rs:121-241
        (x0, y0, z0), (x1, y1, z1) = self.aabbs[self.nodes[oi]["aabb"]]
        xm, ym, zm = x0 + (x1 - x0) / 2.0, y0 + (y1 - y0) / 2.0, z0 + (z1 - z0) / 2.0
        boxes = [((x0, y0, z0), (xm, ym, zm)), ((x0, y0, zm), (xm, ym, z1)), ((xm, y0, zm), (x1, ym, z1)),
                 ((xm, y0, z0), (x1, ym, zm)), ((x0, ym, z0), (xm, y1, zm)), ((x0, ym, zm), (xm, y1, z1)),
                 ((xm, ym, zm), (x1, y1, z1)), ((xm, ym, z0), (x1, y1, zm))]
        out = []
        self.nodes[oi]["kids"] = []
        for b in boxes:
            self.aabbs.append(b)
            self.nodes.append(dict(aabb=len(self.aabbs) - 1, tris=[], kids=[], count=0))
            self.nodes[oi]["kids"].append(len(self.nodes) - 1)
            out.append(len(self.nodes) - 1)
        return out


def intersect_aabb(o, d, box):  # ray.rs:21-54
    (mn, mx) = box
    t1, t2 = _div(mn[0] - o[0], d[0]), _div(mx[0] - o[0], d[0])
    t3, t4 = _div(mn[1] - o[1], d[1]), _div(mx[1] - o[1], d[1])
    t5, t6 = _div(mn[2] - o[2], d[2]), _div(mx[2] - o[2], d[2])
    tmin = _fmax(_fmax(_fmin(t1, t2), _fmin(t3, t4)), _fmin(t5, t6))
    tmax = _fmin(_fmin(_fmax(t1, t2), _fmax(t3, t4)), _fmax(t5, t6))
    if tmax < 0.0 or tmin > tmax:
        return None
    return tmax if tmin < 0.0 else tmin


def intersect_triangle(o, d, tri):  # ray.rs:56-94
    v1, v2, v3 = tri
    e1, e2 = sub(v2, v1), sub(v3, v1)
    h = cross(d, e2)
    a = dot(e1, h)
    if -EPS < a < EPS:
        return None
    f = _div(1.0, a)
    s = sub(o, v1)
    u = f * dot(s, h)
    if u < 0.0 or u > 1.0:
        return None
    q = cross(s, e1)
    v = f * dot(d, q)
    if v < 0.0 or u + v > 1.0:
        return None
    t = f * dot(e2, q)
    return (t, u, v) if t > EPS else None


def intersect_octant(oc, o, d, oi=0, max_t=math.inf):  # ray.rs:104-168, recursive as written
    node = oc.nodes[oi]
    if node["count"] == 0:
        return None
    own, closest = None, max_t
    for ti in node["tris"]:
        r = intersect_triangle(o, d, oc.tris[ti])
        if r is not None and r[0] < closest:
            closest, own = r[0], (r[0], r[1], r[2], ti)
    kids = []
    for c in node["kids"]:
        k = intersect_aabb(o, d, oc.aabbs[oc.nodes[c]["aabb"]])
        if k is not None:
            kids.append((k, c))
    kids.sort(key=lambda kc: kc[0])  # Python's sort is stable, like slice::sort_by (ray.rs:147)
    child, child_t = None, math.inf
    for _, c in kids:
        r = intersect_octant(oc, o, d, c, math.inf)
        if r is not None:
            child, child_t = r, r[0]
            break
    return child if child_t < closest else own


def as_usize(x):  # Rust `as usize`: saturating, NaN -> 0
    if not (x > 0.0):
        return 0
    if x >= 18446744073709551616.0:
        return 2 ** 64 - 1
    return int(x)


def clamp_u8(x):  # f64::clamp(0,255) as u8
    if x != x:
        return 0
    return int(min(max(x, 0.0), 255.0))


class PyTracer:
    """raytracer.rs:22-304 over plain Python data: scene_data = rrt_b200.SceneData."""

    def __init__(self, sd, lights, max_depth=5, offset=1e-4):
        self.sd, self.lights, self.max_depth, self.offset = sd, lights, max_depth, offset
        self.oc = Octree(sd.root_min, sd.root_max)
        for r in sd.v.tolist():
            self.oc.push(tuple(r[0:3]), tuple(r[3:6]), tuple(r[6:9]))

    def _lit(self, p, n, target):  # raytracer.rs:164-188
        d = sub(target, p)
        return intersect_octant(self.oc, add(p, mul(n, self.offset)), d, 0, length(d)) is None

    def _light(self, I, m, inten, n, v, l):  # raytracer.rs:260-304
        ndl = dot(n, l)
        diff = (0.0, 0.0, 0.0) if ndl <= 0.0 else divs(mul(mul(tuple(m.kd), inten), ndl), length(n) * length(l))
        I = add(I, diff)
        spec = (0.0, 0.0, 0.0)
        if m.ns != -1.0:
            r = sub(mul(mul(n, 2.0), dot(n, l)), l)
            rdv = dot(r, v)
            if rdv > 0.0:
                spec = mul(mul(tuple(m.ks), inten), math.pow(_div(rdv, length(r) * length(v)), m.ns))
        return add(I, spec)

    def colour(self, o, d, depth=0):  # raytracer.rs:33-112 -> ((r,g,b), depth-0 hit id)
        h = intersect_octant(self.oc, o, d)
        if h is None:
            return (255, 255, 255), -1
        t, u, v, ti = h
        sd = self.sd
        m = sd.materials[int(sd.material_id[ti])]
        p = add(o, mul(d, t))
        w = 1.0 - u - v
        vt, vn = sd.vt[ti].tolist(), sd.vn[ti].tolist()
        tex = sd.textures[m.tex_id]
        th, tw = tex.shape[0], tex.shape[1]
        tex_x = vt[2] * u + vt[4] * v + vt[0] * w
        tex_y = vt[3] * u + vt[5] * v + vt[1] * w
        tx, ty = as_usize(tex_x * float(tw)) % tw, as_usize(tex_y * float(th)) % th
        col = tex.reshape(-1, 3)[tw * ty + tx].tolist()
        n = add(add(mul(tuple(vn[3:6]), u), mul(tuple(vn[6:9]), v)), mul(tuple(vn[0:3]), w))
        if m.bump_id >= 0:  # raytracer.rs:126-159
            bt = sd.textures[m.bump_id]
            b = bt.reshape(-1, 3)[bt.shape[1] * ty + tx].tolist()
            bv = normalised((float(b[0]), float(b[1]), float(b[2])))
            bv = sub(mul(bv, 2.0), (1.0, 1.0, 1.0))
            tg = cross(n, (0.0, 1.0, 0.0))
            if length(tg) == 0.0:
                tg = cross(n, (0.0, 0.0, 1.0))
            tg = normalised(tg)
            bb = normalised(cross(n, tg))
            n = (dot(bv, tg), dot(bv, bb), dot(bv, n))
        n = normalised(n)
        vdir = (-d[0], -d[1], -d[2])
        I = (0.0, 0.0, 0.0)
        for L in self.lights:  # raytracer.rs:205-255
            if L.kind == 0:
                I = add(I, mul(tuple(m.ka), L.intensity))
            elif L.kind == 2:
                I = self._light(I, m, L.intensity, n, vdir, L.vec)
            else:
                if not self._lit(p, n, L.vec):
                    break
                I = self._light(I, m, L.intensity, n, vdir, sub(L.vec, p))
        local = (col[0] * I[0], col[1] * I[1], col[2] * I[2])
        if m.kr > 0.0 and depth < self.max_depth:  # raytracer.rs:76-102
            ddn = dot(d, n)
            rd = normalised(sub(d, mul(mul(n, 2.0), ddn)))
            rc, _ = self.colour(add(p, mul(n, self.offset)), rd, depth + 1)
            fin = add(mul(local, 1.0 - m.kr), mul((float(rc[0]), float(rc[1]), float(rc[2])), m.kr))
            return (clamp_u8(fin[0]), clamp_u8(fin[1]), clamp_u8(fin[2])), ti
        return (clamp_u8(local[0]), clamp_u8(local[1]), clamp_u8(local[2])), ti
