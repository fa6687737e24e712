"""Deterministic ray sets of the bit-exactness stress test (tests/test_gpu_stress.py) — shared by the fixture generator
tools/make_ray_stress.py (which traces them with the reference, oracle/_ref) and the GPU test (which traces them through the
C ABI). Nothing here needs the oracle or a GPU; the scene's extent comes from the host loader (qg_host_scene_load_xml).

Per scene: N closest-hit rays in three classes, then N any-hit rays from the hit points of the first two classes.
  A  outside-in      origin uniform in ball(c, 1.5 R), aimed at a uniform point of ball(c, 0.5 R)
  B  uniform         origin uniform in ball(c, 1.5 R), direction uniform on the sphere
  C  from surfaces   origin = hit point of an A/B ray (o + z d in fp32, z being the distance both implementations must agree
                     on bit for bit); direction uniform / toward another hit point / toward another hit point of the SAME
                     scene node (a ray along the wall or across the mesh it starts on: grazing and coplanar cases)
  S  shadow          origin = hit point, toward a uniform point of ball(c, R) with tmax = its distance (3/4) or uniform
                     direction with tmax = 1e30 (1/4)
numpy's PCG64 stream is stable across numpy versions (NEP 19)."""
import ctypes as C
import hashlib
import os

import numpy as np

SEED = 0x5EED
BLOCK = 4096
BIG = np.float32(1.0e30)


def scene_sphere(xml_path, asset_dir):
    """centre and radius of the scene: world-space bounds of the sphere and mesh nodes' boxes (of the planes when there is nothing else)"""
    from qaray_b200 import capi
    lib = capi.load()
    hs = C.c_void_p()
    capi.check(lib, lib.qg_host_scene_load_xml(os.fsencode(xml_path), os.fsencode(asset_dir), C.byref(hs)))
    d = lib.qg_host_scene_desc(hs).contents
    lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
    solid = any(d.nodes[i].object_type in (1, 3) for i in range(d.num_nodes))  # spheres / meshes: the walls (planes) only frame them
    for i in range(d.num_nodes):
        nd = d.nodes[i]
        if nd.object_type == 0 or (solid and nd.object_type == 2):
            continue
        if nd.object_type == 3:
            m = d.meshes[nd.mesh]
            bmin, bmax = np.array(m.bound_min[:], np.float64), np.array(m.bound_max[:], np.float64)
        else:
            bmin, bmax = -np.ones(3), np.ones(3)
        corners = np.array([[x, y, z] for x in (bmin[0], bmax[0]) for y in (bmin[1], bmax[1]) for z in (bmin[2], bmax[2])])
        j = i
        while j >= 0:  # FromNodeCoords up the chain: p = tm * p + pos (src/core/node.cpp:127-139)
            xf = d.nodes[j].xf
            tm = np.array(xf.tm[:], np.float64).reshape(3, 3).T  # column-major
            corners = corners @ tm.T + np.array(xf.pos[:], np.float64)
            j = d.nodes[j].parent
        lo, hi = np.minimum(lo, corners.min(0)), np.maximum(hi, corners.max(0))
    lib.qg_host_scene_destroy(hs)
    c = 0.5 * (lo + hi)
    r = 0.5 * float(np.linalg.norm(hi - lo))
    return c.astype(np.float32), np.float32(min(r, 200.0))


def _ball(rng, n, c, r):
    v = rng.normal(size=(n, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    return (c + v * (r * rng.random(n) ** (1.0 / 3.0))[:, None]).astype(np.float32)


def _sphere(rng, n):
    v = rng.normal(size=(n, 3))
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(np.float32)


def _norm(v):
    v = v.astype(np.float32)
    n = np.sqrt((v * v).sum(1, dtype=np.float32), dtype=np.float32)
    n[n == 0] = 1
    return (v / n[:, None]).astype(np.float32)


def rays_ab(c, r, n):
    """classes A and B: (n, 6) float32 = origin, direction"""
    rng = np.random.default_rng(SEED)
    h = n // 2
    oa, ta = _ball(rng, h, c, 1.5 * r), _ball(rng, h, c, 0.5 * r)
    ob, db = _ball(rng, n - h, c, 1.5 * r), _sphere(rng, n - h)
    return np.concatenate([np.concatenate([oa, _norm(ta - oa)], 1), np.concatenate([ob, db], 1)], 0)


def hit_points(rays, hit, z):
    m = hit == 1
    return (rays[m, :3] + rays[m, 3:6] * z[m].astype(np.float32)[:, None]).astype(np.float32), np.nonzero(m)[0]


def rays_c(rays_ab_, hit, node, z, n):
    """class C from the A/B results (hit flags, node ids and distances are what the digest of A/B pins)"""
    rng = np.random.default_rng(SEED + 1)
    p, src = hit_points(rays_ab_, hit, z)
    if len(p) == 0:
        return np.zeros((0, 6), np.float32)
    pick = rng.integers(0, len(p), n)
    o = p[pick]
    d = _sphere(rng, n)
    q = n // 4
    other = p[rng.integers(0, len(p), q)]
    d[q:2 * q] = _norm(other - o[q:2 * q])
    # toward another hit point of the same scene node: sort the hits by node, take the neighbour in that order
    order = np.argsort(node[src], kind="stable")
    rank = np.empty(len(p), np.int64)
    rank[order] = np.arange(len(p))
    nb = order[(rank[pick[2 * q:3 * q]] + 1 + rng.integers(0, 8, q)) % len(p)]
    d[2 * q:3 * q] = _norm(p[nb] - o[2 * q:3 * q])
    return np.concatenate([o, d], 1).astype(np.float32)


def rays_s(rays_ab_, hit, z, c, r, n):
    """any-hit rays: (n, 6) rays and (n,) tmax"""
    rng = np.random.default_rng(SEED + 2)
    p, _ = hit_points(rays_ab_, hit, z)
    if len(p) == 0:
        return np.zeros((0, 6), np.float32), np.zeros(0, np.float32)
    o = p[rng.integers(0, len(p), n)]
    tgt = _ball(rng, n, c, r)
    v = (tgt - o).astype(np.float32)
    tmax = np.sqrt((v * v).sum(1, dtype=np.float32), dtype=np.float32)
    d = _norm(v)
    q = n // 4
    d[:q] = _sphere(rng, q)
    tmax[:q] = BIG
    return np.concatenate([o, d], 1).astype(np.float32), tmax.astype(np.float32)


def pack_closest(hit, node, face, z):
    """the record the digests cover: hit, node, face id, distance bits; canonical on a miss"""
    rec = np.zeros(len(hit), np.dtype([("hit", "u1"), ("node", "<i4"), ("face", "<i4"), ("z", "<u4")]))
    m = hit == 1
    rec["hit"] = m
    rec["node"] = np.where(m, node, -1)
    rec["face"] = np.where(m, face, -1)
    rec["z"] = np.where(m, np.asarray(z, np.float32).view(np.uint32), 0)
    return rec


def digests(rec):
    """SHA-256 of every BLOCK records -> (nblocks, 32) uint8"""
    raw = np.ascontiguousarray(rec).view(np.uint8).reshape(len(rec), -1)
    out = []
    for b in range(0, len(rec), BLOCK):
        out.append(np.frombuffer(hashlib.sha256(raw[b:b + BLOCK].tobytes()).digest(), np.uint8))
    return np.array(out, np.uint8).reshape(-1, 32)
