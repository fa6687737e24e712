"""Bit-exactness under stress: 2^20 closest-hit and 2^20 any-hit rays per mesh scene through qg_trace_closest / qg_trace_any
against the reference's results (tests/golden/rays_<scene>.npz, tools/make_ray_stress.py: SHA-256 digests per 4096 rays of
hit / node / face id / distance bits, every 64th record in full). The ray sets (tests/raystress.py) include rays that start
ON surfaces and run along them — what the traversal's two work-saving culls (degenerate axis, boxes behind the origin,
qaray_b200/csrc/traverse.cuh) must not get wrong. Match: src/scene/scene.cpp:35-74, src/objects/objects.cpp:212-420."""
import os

import numpy as np
import pytest

import raystress as rs
from conftest import ASSETS, GOLDEN_DIR, scene_xml
from qaray_b200 import capi
from qaray_b200.renderer import Scene

pytestmark = pytest.mark.gpu


def stress_names():
    if not os.path.isdir(GOLDEN_DIR):
        return []
    return sorted(f[5:-4] for f in os.listdir(GOLDEN_DIR) if f.startswith("rays_") and f.endswith(".npz"))


def to_struct(r6):
    r = np.zeros(len(r6), capi.RAY_DTYPE)
    r["p"] = r["xp"] = r["yp"] = r6[:, :3]
    r["dir"] = r["xdir"] = r["ydir"] = r6[:, 3:6]
    return r


def compare(what, rec, dig, sub):
    """digest per block; on a mismatch name the block and, where the stored sample allows, the first differing record"""
    mine = rs.digests(rec)
    assert mine.shape == dig.shape, "%s: %d blocks, fixture has %d" % (what, len(mine), len(dig))
    bad = np.nonzero((mine != dig).any(axis=1))[0]
    if len(bad) == 0:
        return
    msub = rec[::64]
    diff = np.nonzero(msub != sub)[0] if msub.dtype.names is None else np.nonzero(
        np.any([msub[f] != sub[f] for f in msub.dtype.names], axis=0))[0]
    detail = ""
    if len(diff):
        k = int(diff[0])
        detail = "; ray %d: got %s, reference %s" % (64 * k, msub[k], sub[k])
    pytest.fail("%s: %d of %d blocks of %d rays differ from the reference (first block %d)%s" % (
        what, len(bad), len(dig), rs.BLOCK, int(bad[0]), detail))


@pytest.mark.parametrize("name", stress_names())
def test_two_million_rays_bit_equal(name):
    g = np.load(os.path.join(GOLDEN_DIR, "rays_%s.npz" % name))
    n = int(g["n"])
    xml = scene_xml(name)
    c, r = rs.scene_sphere(xml, ASSETS)
    assert np.array_equal(c, g["centre"]) and np.float32(r) == g["radius"], "the host loader's scene bounds moved"
    sc = Scene.load(xml, ASSETS)
    ab = rs.rays_ab(c, r, n // 2)
    hab = sc.trace_closest(to_struct(ab))
    compare("closest hit, classes A/B", rs.pack_closest(hab["hit"], hab["node"], hab["face"], hab["z"]), g["dig_ab"], g["sub_ab"])
    cc = rs.rays_c(ab, hab["hit"], hab["node"], hab["z"], n // 2)
    hc = sc.trace_closest(to_struct(cc))
    compare("closest hit, class C (rays from surfaces)", rs.pack_closest(hc["hit"], hc["node"], hc["face"], hc["z"]), g["dig_c"], g["sub_c"])
    s6, tmax = rs.rays_s(ab, hab["hit"], hab["z"], c, r, n)
    occ = sc.trace_any(s6, tmax)
    compare("any hit", occ.astype(np.uint8), g["dig_s"], g["sub_s"])
    sc.close()
    assert int(hab["hit"].sum()) == int(g["hits"][0]) and int(hc["hit"].sum()) == int(g["hits"][1]) and int(occ.sum()) == int(g["hits"][2])
