#!/usr/bin/env python3
"""Deterministic stand-in assets for the reference's scene inputs.

The reference's CMake deploys assets/examples/*.obj|mtl|png and assets/trc2017/* (CMakeLists.txt:31-44), but
those files are not in the reference tree (only assets/renderings/*.png exist). Four of the five BASELINE
configs need them. This script writes closed, consistently wound, per-vertex-normal meshes and procedural
RGB8 textures under the SAME file names, sized for the role each plays in the XMLs, from a fixed seed.
Both the CPU oracle and the CUDA path read these same files, so parity is unaffected; every throughput
number measured on them is labelled "stand-in geometry".

    python tools/gen_assets.py [outdir=assets] [--scale S]     (S<1 shrinks triangle counts for quick tests)
"""
import os
import sys
import zlib
import struct
import numpy as np

SEED = 0x5EED


# ------------------------------------------------------------------------------------------ mesh helpers
class Mesh:
    def __init__(self, v, f, vt=None, mtl=None):
        self.v = np.asarray(v, np.float64).reshape(-1, 3)
        self.f = np.asarray(f, np.int64).reshape(-1, 3)
        self.vt = None if vt is None else np.asarray(vt, np.float64).reshape(-1, 2)
        self.mtl = mtl  # per-face material index array or None

    def transformed(self, scale=1.0, rot=None, offset=(0, 0, 0)):
        v = self.v * np.asarray(scale, np.float64)
        if rot is not None:
            v = v @ np.asarray(rot, np.float64).T
        return Mesh(v + np.asarray(offset, np.float64), self.f.copy(), self.vt, self.mtl)


def merge(meshes, mtl_ids=None):
    vs, fs, vts, ms = [], [], [], []
    off = 0
    have_vt = all(m.vt is not None for m in meshes)
    for k, m in enumerate(meshes):
        vs.append(m.v)
        fs.append(m.f + off)
        if have_vt:
            vts.append(m.vt)
        if mtl_ids is not None:
            ms.append(np.full(len(m.f), mtl_ids[k], np.int64))
        off += len(m.v)
    return Mesh(np.concatenate(vs), np.concatenate(fs), np.concatenate(vts) if have_vt else None,
                np.concatenate(ms) if mtl_ids is not None else None)


def vertex_normals(m):
    a, b, c = m.v[m.f[:, 0]], m.v[m.f[:, 1]], m.v[m.f[:, 2]]
    fn = np.cross(b - a, c - a)
    vn = np.zeros_like(m.v)
    for k in range(3):
        np.add.at(vn, m.f[:, k], fn)
    ln = np.linalg.norm(vn, axis=1, keepdims=True)
    ln[ln == 0] = 1.0
    return vn / ln


def grid_faces(nu, nv, wrap_u):
    """faces of a (nv rows) x (nu columns) vertex grid, row-major; CCW when u goes right and v goes up seen from outside"""
    cols = nu if wrap_u else nu - 1
    i, j = np.meshgrid(np.arange(cols), np.arange(nv - 1))
    i, j = i.ravel(), j.ravel()
    i1 = (i + 1) % nu
    a = j * nu + i
    b = j * nu + i1
    c = (j + 1) * nu + i1
    d = (j + 1) * nu + i
    return np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)])


def lathe(radius_fn, z0, z1, n_theta, n_z, axis="z"):
    """closed surface of revolution r = radius_fn(theta, t) for t in [0,1] along the axis, capped at both ends"""
    th = np.arange(n_theta) / n_theta * 2 * np.pi
    t = np.linspace(0.0, 1.0, n_z)
    T, TH = np.meshgrid(t, th, indexing="ij")
    R = radius_fn(TH, T)
    Z = z0 + (z1 - z0) * T
    v = np.stack([R * np.cos(TH), R * np.sin(TH), Z], -1).reshape(-1, 3)
    vt = np.stack([TH / (2 * np.pi), T], -1).reshape(-1, 2)
    f = grid_faces(n_theta, n_z, True)
    # caps: fans around axis points
    nb = len(v)
    v = np.concatenate([v, [[0, 0, z0], [0, 0, z1]]])
    vt = np.concatenate([vt, [[0.5, 0.0], [0.5, 1.0]]])
    i = np.arange(n_theta)
    i1 = (i + 1) % n_theta
    bot = np.stack([np.full(n_theta, nb), i1, i], 1)
    top0 = (n_z - 1) * n_theta
    top = np.stack([np.full(n_theta, nb + 1), top0 + i, top0 + i1], 1)
    f = np.concatenate([f, bot, top])
    m = Mesh(v, f, vt)
    if axis == "y":  # (x, y, z) -> (x, z, -y): rotate so the lathe axis is +y
        m = Mesh(np.stack([m.v[:, 0], m.v[:, 2], -m.v[:, 1]], 1), m.f, m.vt)
    return m


def tube(path_fn, radius_fn, n_s, n_c):
    """closed tube around a planar (x,z) path, used for handles / spouts"""
    s = np.linspace(0.0, 1.0, n_s)
    P = path_fn(s)                       # (n_s, 3)
    dP = np.gradient(P, axis=0)
    Tn = dP / np.linalg.norm(dP, axis=1, keepdims=True)
    B = np.tile(np.array([0.0, 1.0, 0.0]), (n_s, 1))
    Nn = np.cross(B, Tn)
    Nn /= np.linalg.norm(Nn, axis=1, keepdims=True)
    ph = np.arange(n_c) / n_c * 2 * np.pi
    r = radius_fn(s)
    v = (P[:, None, :] + r[:, None, None] * (np.cos(ph)[None, :, None] * Nn[:, None, :] +
                                             np.sin(ph)[None, :, None] * B[:, None, :])).reshape(-1, 3)
    vt = np.stack(np.meshgrid(ph / (2 * np.pi), s), -1).reshape(-1, 2)
    f = grid_faces(n_c, n_s, True)
    nb = len(v)
    v = np.concatenate([v, [P[0], P[-1]]])
    vt = np.concatenate([vt, [[0.5, 0.0], [0.5, 1.0]]])
    i = np.arange(n_c)
    i1 = (i + 1) % n_c
    f = np.concatenate([f, np.stack([np.full(n_c, nb), i1, i], 1),
                        np.stack([np.full(n_c, nb + 1), (n_s - 1) * n_c + i, (n_s - 1) * n_c + i1], 1)])
    return Mesh(v, f, vt)


def box(lo, hi, n=1):
    """axis-aligned box, each face an n x n grid (2 n^2 triangles), outward winding"""
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    parts = []
    g = np.linspace(0.0, 1.0, n + 1)
    U, V = np.meshgrid(g, g)
    U, V = U.ravel(), V.ravel()
    f = grid_faces(n + 1, n + 1, False)
    for ax in range(3):
        for side in (0, 1):
            a1, a2 = (ax + 1) % 3, (ax + 2) % 3
            v = np.zeros((len(U), 3))
            v[:, ax] = hi[ax] if side else lo[ax]
            if side:
                v[:, a1] = lo[a1] + (hi[a1] - lo[a1]) * U
                v[:, a2] = lo[a2] + (hi[a2] - lo[a2]) * V
            else:
                v[:, a1] = lo[a1] + (hi[a1] - lo[a1]) * V
                v[:, a2] = lo[a2] + (hi[a2] - lo[a2]) * U
            parts.append(Mesh(v, f, np.stack([U, V], 1)))
    return merge(parts)


def icosphere(sub):
    t = (1 + 5 ** 0.5) / 2
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2],
                  [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5],
                  [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], np.int64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    for _ in range(sub):
        cache = {}
        vl = list(map(tuple, v))
        nf = []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                p = (np.array(vl[a]) + np.array(vl[b])) / 2
                vl.append(tuple(p / np.linalg.norm(p)))
                cache[key] = len(vl) - 1
            return cache[key]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        v, f = np.array(vl), np.array(nf, np.int64)
    return Mesh(v, f)


def smoothstep(x):
    x = np.clip(x, 0, 1)
    return x * x * (3 - 2 * x)


# ------------------------------------------------------------------------------------------ writers
def write_obj(path, m, with_normals=True, mtllib=None, mtl_names=None, group=None):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    vn = vertex_normals(m) if with_normals else None
    out = ["# stand-in asset generated by tools/gen_assets.py (seed 0x5EED)"]
    if mtllib:
        out.append("mtllib %s" % mtllib)
    if group:
        out.append("g %s" % group)
    out += ["v %.6f %.6f %.6f" % tuple(p) for p in m.v]
    if m.vt is not None:
        out += ["vt %.6f %.6f" % tuple(p) for p in m.vt]
    if vn is not None:
        out += ["vn %.6f %.6f %.6f" % tuple(p) for p in vn]
    f1 = m.f + 1
    if m.vt is not None and vn is not None:
        fmt = "f %d/%d/%d %d/%d/%d %d/%d/%d"
        rows = np.repeat(f1, 3, axis=1)
    elif vn is not None:
        fmt = "f %d//%d %d//%d %d//%d"
        rows = np.repeat(f1, 2, axis=1)
    elif m.vt is not None:
        fmt = "f %d/%d %d/%d %d/%d"
        rows = np.repeat(f1, 2, axis=1)
    else:
        fmt = "f %d %d %d"
        rows = f1
    if m.mtl is None:
        out += [fmt % tuple(r) for r in rows]
    else:
        cur = None
        for r, k in zip(rows, m.mtl):
            if k != cur:
                out.append("usemtl %s" % mtl_names[k])
                cur = k
            out.append(fmt % tuple(r))
    with open(path, "w") as fp:
        fp.write("\n".join(out) + "\n")
    return len(m.f)


def write_png(path, rgb):
    """minimal RGB8 PNG writer (zlib level 6, filter 0) so the bytes do not depend on PIL"""
    os.makedirs(os.path.dirname(path), exist_ok=True)
    rgb = np.ascontiguousarray(rgb, np.uint8)
    h, w, _ = rgb.shape
    raw = b"".join(b"\x00" + rgb[y].tobytes() for y in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)
    png = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) + \
        chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b"")
    with open(path, "wb") as fp:
        fp.write(png)


# ------------------------------------------------------------------------------------------ models
def teapot(n_theta, n_z, n_tube):
    """teapot-like closed body (z-up, base at z=0, ~15.7 tall, radius ~10) + spout + handle, all closed solids"""
    def body(th, t):
        prof = 7.5 + 2.8 * np.sin(np.pi * np.clip(t * 0.92, 0, 1)) ** 0.8 - 1.8 * smoothstep((t - 0.78) / 0.22) * 2.4
        lid = 2.2 * np.exp(-((t - 0.97) / 0.03) ** 2)
        return np.maximum(prof + lid, 0.6)
    b = lathe(body, 0.0, 15.75, n_theta, n_z)

    def spout_path(s):
        return np.stack([9.0 + 8.5 * s, 0 * s, 5.0 + 7.5 * s ** 1.6], 1)
    sp = tube(spout_path, lambda s: 2.2 - 1.3 * s, max(8, n_tube), max(8, n_tube // 2))

    def handle_path(s):
        a = -0.5 * np.pi + s * np.pi
        return np.stack([-9.2 - 5.5 * np.cos(a), 0 * s, 8.0 + 4.6 * np.sin(a)], 1)
    hd = tube(handle_path, lambda s: 1.0 + 0 * s, max(8, n_tube), max(8, n_tube // 2))
    return merge([b, sp, hd])


def doll(n_theta, n_z):
    """y-up figurine, height ~1 (the XMLs scale it by 20)"""
    def prof(th, t):
        body = 0.16 * np.sin(np.pi * np.clip(t / 0.62, 0, 1)) ** 0.7 + 0.05
        neck = 0.05
        head = 0.13 * np.sqrt(np.clip(1 - ((t - 0.84) / 0.15) ** 2, 0, 1))
        r = np.where(t < 0.62, body, np.where(t < 0.70, neck, head))
        return np.maximum(r, 0.01) * (1 + 0.04 * np.cos(4 * th))
    return lathe(prof, 0.0, 1.0, n_theta, n_z, axis="y")


def suzanne(sub):
    m = icosphere(sub)
    v = m.v.copy()
    x, y, z = v[:, 0], v[:, 1], v[:, 2]
    r = 1.0 + 0.25 * np.exp(-((x - 0.9) ** 2 + (y - 0.3) ** 2) * 6) + 0.25 * np.exp(-((x + 0.9) ** 2 + (y - 0.3) ** 2) * 6) \
        + 0.18 * np.exp(-((y + 0.5) ** 2 + x ** 2) * 5) * (z > 0) - 0.12 * np.exp(-((x - 0.35) ** 2 + (y - 0.25) ** 2) * 30) * (z > 0) \
        - 0.12 * np.exp(-((x + 0.35) ** 2 + (y - 0.25) ** 2) * 30) * (z > 0)
    v = v * r[:, None] * np.array([1.0, 0.85, 0.8])
    return Mesh(v, m.f)


def tower(n_theta, n_z, height, base_r, windows, crenel):
    """y-up tower: displaced lathe with window recesses and battlements"""
    def prof(th, t):
        r = base_r * (1.0 - 0.35 * t)
        for k, (t0, t1, n) in enumerate(windows):
            inz = smoothstep((t - t0) / 0.01) * (1 - smoothstep((t - t1) / 0.01))
            inth = (np.cos(n * th + k) > 0.55).astype(np.float64)
            r = r - 0.18 * base_r * inz * inth
        r = r + 0.22 * base_r * smoothstep((t - 0.90) / 0.01) * ((np.cos(crenel * th) > 0) * 0.5 + 0.5)
        r = r + 0.06 * base_r * (np.sin(40 * np.pi * t) > 0.8) * (t < 0.88)
        return r
    return lathe(prof, 0.0, height, n_theta, n_z, axis="y")


def church(n_theta, n_z):
    nave = box([-1.6, 0.0, -1.0], [1.6, 1.6, 1.0], n=max(2, n_z // 8))
    def steeple(th, t):
        sq = 0.55 / np.maximum(np.abs(np.cos(th)), np.abs(np.sin(th)))   # square cross-section
        r = np.where(t < 0.55, sq, sq * (1 - (t - 0.55) / 0.45) + 0.01)
        r = r - 0.12 * (np.cos(4 * th) > 0.7) * (np.abs(t - 0.4) < 0.08)
        return r
    st = lathe(steeple, 0.0, 4.2, n_theta, n_z, axis="y").transformed(offset=(-1.0, 0.0, 0.0))
    def roof(th, t):
        return 1.25 * (1 - t) / np.maximum(np.abs(np.cos(th)), 0.6 * np.abs(np.sin(th))) * 0.8 + 0.01
    rf = lathe(roof, 1.6, 2.6, max(16, n_theta // 4), max(8, n_z // 8), axis="y").transformed(scale=(1.2, 1.0, 1.0), offset=(0.6, 0, 0))
    return merge([nave, st, rf])


def rot_x(deg):
    a = np.radians(deg)
    return np.array([[1, 0, 0], [0, np.cos(a), -np.sin(a)], [0, np.sin(a), np.cos(a)]])


def xmas_parts(q):
    """z-up parts of the snow-globe box of trc_scene_xmas.xml (group scale 0.61)"""
    rng = np.random.RandomState(SEED)
    n = max(2, int(8 * q))
    parts = {}
    parts["Box"] = merge([box([-8, -5, 0], [8, 5, 0.6], n), box([-8, -5, 0.6], [-7.6, 5, 9], n), box([7.6, -5, 0.6], [8, 5, 9], n),
                          box([-8, 4.6, 0.6], [8, 5, 9], n)])
    parts["BoxBugFix"] = box([-8, -5, 9], [8, 5, 9.4], n)
    houses, mids = [], []
    for k in range(7):
        x = -6.2 + k * 2.05
        h = 2.2 + 1.6 * rng.rand()
        d = 1.2 + 0.8 * rng.rand()
        houses.append(box([x, 1.0, 0.6], [x + 1.7, 1.0 + d, 0.6 + h], n)); mids.append(k % 3)
        # gable roof as a 4-sided squashed pyramid (closed)
        def roof(th, t, w=1.0):
            return (1 - t) * 1.25 / np.maximum(np.abs(np.cos(th)), np.abs(np.sin(th))) + 0.01
        r = lathe(roof, 0.0, 1.1, max(8, 4 * n), max(4, n))
        houses.append(r.transformed(scale=(0.85, d / 2 * 0.95, 1.0), offset=(x + 0.85, 1.0 + d / 2, 0.6 + h))); mids.append(3)
    parts["Houses"] = merge(houses, mids)
    panes_l, panes_d, lamps, stars = [], [], [], []
    for k in range(7):
        x = -6.2 + k * 2.05
        for w in range(2):
            p = box([x + 0.25 + 0.8 * w, 0.93, 1.3], [x + 0.65 + 0.8 * w, 0.99, 2.0], 1)
            (panes_l if (k + w) % 2 == 0 else panes_d).append(p)
    for k in range(5):
        x = -6.5 + k * 3.2
        lamps.append(box([x - 0.05, -1.0, 0.6], [x + 0.05, -0.9, 3.0], 1))
        s = icosphere(1 if q < 0.5 else 2)
        lamps.append(s.transformed(scale=0.22, offset=(x, -0.95, 3.15)))
    for k in range(24):
        s = icosphere(0 if q < 0.5 else 1)
        stars.append(s.transformed(scale=0.07 + 0.05 * rng.rand(), offset=(-7 + 14 * rng.rand(), -3 + 6 * rng.rand(), 6.5 + 2.2 * rng.rand())))
    parts["LightedGlass"] = merge(panes_l)
    parts["DarkGlass"] = merge(panes_d)
    parts["StreetLights"] = merge(lamps)
    parts["Stars"] = merge(stars)
    return parts


def cornell(n):
    """cornell-box-like room with two blocks and a .mtl (z up is y here, like the classic file)"""
    wall = []
    mids = []
    L = 550.0
    room = box([0, 0, 0], [L, L, L], n)
    # flip winding so that the room faces inward
    room = Mesh(room.v, room.f[:, ::-1], room.vt)
    # split by face: box() emits 6 faces of 2 n^2 triangles: -x,+x,-y,+y,-z,+z
    per = 2 * n * n
    mtl = np.zeros(len(room.f), np.int64)
    mtl[0 * per:1 * per] = 1   # -x red
    mtl[1 * per:2 * per] = 2   # +x green
    keep = np.ones(len(room.f), bool)
    keep[5 * per:6 * per] = False   # open front (+z)
    room = Mesh(room.v, room.f[keep], room.vt, mtl[keep])
    short = box([130, 0, 65], [295, 165, 230], max(1, n // 2))
    tall = box([265, 0, 295], [430, 330, 460], max(1, n // 2))
    light = box([213, 548.0, 227], [343, 548.7, 332], 1)
    parts = [room, Mesh(short.v, short.f, short.vt, np.zeros(len(short.f), np.int64)),
             Mesh(tall.v, tall.f, tall.vt, np.zeros(len(tall.f), np.int64)),
             Mesh(light.v, light.f, light.vt, np.full(len(light.f), 3, np.int64))]
    m = merge(parts, None)
    m.mtl = np.concatenate([p.mtl for p in parts])   # per-face material ids survive the merge (usemtl lines)
    return m, None


# ------------------------------------------------------------------------------------------ textures
def value_noise(n, cells, rng):
    g = rng.rand(cells + 1, cells + 1)
    g[-1, :] = g[0, :]
    g[:, -1] = g[:, 0]
    x = np.arange(n) / n * cells
    i = x.astype(int)
    f = x - i
    f = f * f * (3 - 2 * f)
    a = g[i][:, i]
    b = g[i][:, i + 1]
    c = g[i + 1][:, i]
    d = g[i + 1][:, i + 1]
    fx, fy = f[None, :], f[:, None]
    return a * (1 - fx) * (1 - fy) + b * fx * (1 - fy) + c * (1 - fx) * fy + d * fx * fy


def fbm(n, rng, octaves=5):
    out = np.zeros((n, n))
    amp, cells, tot = 1.0, 4, 0.0
    for _ in range(octaves):
        out += amp * value_noise(n, cells, rng)
        tot += amp
        amp *= 0.5
        cells *= 2
    return out / tot


def textures(outdir, n=512):
    rng = np.random.RandomState(SEED)
    c = fbm(n, rng)
    c = smoothstep((c - 0.42) / 0.25)
    sky = np.stack([0.35 + 0.65 * c, 0.55 + 0.45 * c, 0.95 + 0.05 * c], -1)
    write_png(os.path.join(outdir, "clouds.png"), (np.clip(sky, 0, 1) * 255 + 0.5).astype(np.uint8))
    y, x = np.mgrid[0:n, 0:n]
    row = y // 32
    bx = (x + (row % 2) * 32) % 64
    mortar = (y % 32 < 3) | (bx < 3)
    nb = fbm(n, rng, 4)
    br = np.stack([0.62 + 0.25 * nb, 0.25 + 0.12 * nb, 0.18 + 0.08 * nb], -1)
    br[mortar] = [0.75, 0.74, 0.70]
    write_png(os.path.join(outdir, "bricks.png"), (np.clip(br, 0, 1) * 255 + 0.5).astype(np.uint8))
    nw = fbm(n, rng, 4)
    rings = 0.5 + 0.5 * np.sin((x / n * 18 + 6 * nw) * 2 * np.pi)
    wd = np.stack([0.55 + 0.25 * rings, 0.35 + 0.18 * rings, 0.16 + 0.10 * rings], -1)
    write_png(os.path.join(outdir, "wood.png"), (np.clip(wd, 0, 1) * 255 + 0.5).astype(np.uint8))


HOUSES_MTL = """# stand-in material library generated by tools/gen_assets.py
newmtl plaster
Kd 0.80 0.76 0.68
Ks 0.10 0.10 0.10
Ns 20
Ni 1.0
illum 2
newmtl brick
Kd 0.55 0.22 0.16
Ks 0.05 0.05 0.05
Ns 10
Ni 1.0
illum 2
newmtl timber
Kd 0.36 0.24 0.12
Ks 0.20 0.20 0.20
Ns 40
Ni 1.0
illum 2
newmtl roof
Kd 0.30 0.05 0.05
Ks 0.60 0.60 0.60
Ns 80
Ni 1.0
illum 3
"""

CORNELL_MTL = """# stand-in material library generated by tools/gen_assets.py
newmtl white
Kd 0.73 0.73 0.73
Ks 0 0 0
Ns 10
illum 2
newmtl red
Kd 0.65 0.05 0.05
Ks 0 0 0
Ns 10
illum 2
newmtl green
Kd 0.12 0.45 0.15
Ks 0 0 0
Ns 10
illum 2
newmtl light
Kd 0.78 0.78 0.78
Ks 0 0 0
Ns 10
illum 2
"""


def generate(outdir, q=1.0, quiet=False):
    """q scales tessellation (1.0 = full-size stand-ins). Returns {relative file: triangle count}."""
    counts = {}

    def s(n, lo=8):
        return max(lo, int(round(n * np.sqrt(q))))

    def emit(rel, m, **kw):
        counts[rel] = write_obj(os.path.join(outdir, rel), m, **kw)
        if not quiet:
            print("  %-28s %8d triangles" % (rel, counts[rel]))

    emit("teapot-high.obj", teapot(s(64), s(40), s(24)))                  # ~6k
    emit("teapot-low.obj", teapot(s(24), s(16), s(10)))                   # ~1k
    emit("wooddoll.obj", doll(s(48), s(52)))                              # ~5k, has uv
    sz = suzanne(4 if q >= 0.5 else 2)
    emit("suzanne.obj", Mesh(sz.v, sz.f))                                 # 5k, no uv
    emit("trc2017/TeapotSolid.obj", teapot(s(180), s(120), s(64)))        # ~50k
    emit("trc2017/tower.obj", tower(s(400), s(250), 4.6, 0.95, [(0.18, 0.26, 6), (0.42, 0.52, 8), (0.66, 0.74, 6)], 12))   # ~200k
    emit("trc2017/church1.obj", church(s(260), s(180)))                   # ~100k
    for name, m in xmas_parts(q).items():
        if name == "Houses":
            emit("trc2017/Houses.obj", m, mtllib="Houses.mtl", mtl_names=["plaster", "brick", "timber", "roof"])
            with open(os.path.join(outdir, "trc2017/Houses.mtl"), "w") as fp:
                fp.write(HOUSES_MTL)
        else:
            emit("trc2017/%s.obj" % name, m)
    cb, _ = cornell(max(1, int(4 * q)))
    emit("examples/cornell_box.obj", cb, mtllib="cornell_box.mtl", mtl_names=["white", "red", "green", "light"])
    with open(os.path.join(outdir, "examples/cornell_box.mtl"), "w") as fp:
        fp.write(CORNELL_MTL)
    textures(outdir, 512 if q >= 0.5 else 128)
    for t in ("clouds.png", "bricks.png"):
        with open(os.path.join(outdir, t), "rb") as a, open(os.path.join(outdir, "examples", t), "wb") as b:
            b.write(a.read())
    emit("examples/wooddoll.obj", doll(s(48), s(52)))
    with open(os.path.join(outdir, "STANDIN_ASSETS.txt"), "w") as fp:
        fp.write("stand-in geometry/textures generated by tools/gen_assets.py q=%g seed=0x5EED\n" % q)
        for k in sorted(counts):
            fp.write("%s %d\n" % (k, counts[k]))
    return counts


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    q = 1.0
    if "--scale" in sys.argv:
        q = float(sys.argv[sys.argv.index("--scale") + 1])
        args = [a for a in args if a != sys.argv[sys.argv.index("--scale") + 1]]
    out = args[0] if args else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "assets")
    print("writing stand-in assets to %s (q=%g)" % (out, q))
    generate(out, q)
