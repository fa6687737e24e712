// trace.cuh — ray/scene intersection device code. MUST be compiled with --fmad=false: every expression below
// keeps the reference's operation order so hit/miss, primitive id and t are bit-identical to its CPU code
// (oracle build: -O2 -ffp-contract=off). Division and sqrt are the IEEE-rounded defaults of nvcc
// (-prec-div=true -prec-sqrt=true), denormals are kept (-ftz=false).
//
//   Scene::TraceNodeNormal / TraceNodeShadow   src/scene/scene.cpp:35-74      -> traverse_pool (traverse.cuh)
//   Node::ToNodeCoords / FromNodeCoords        src/core/node.cpp:112-139      -> to_node / from_node
//   Sphere::IntersectRay                       src/objects/objects.cpp:55-141
//   Plane::IntersectRay                        src/objects/objects.cpp:149-208
//   TriObj::IntersectRay + Box::IntersectRay   src/objects/objects.cpp:310-322, src/core/box.cpp:94-128
//   TriObj::TraceBVHNode                       src/objects/objects.cpp:324-420
//   TriObj::IntersectTriangle                  src/objects/objects.cpp:212-306
#pragma once
#include "dev_scene.h"

namespace qb {

struct Ray3 {
  V3 p, d;
};

struct TraceCounters {
  unsigned long long nodes, tris, objects;
};

#ifdef QB_COUNT
#define QB_CNT(field) (cnt.field++)
#else
#define QB_CNT(field) ((void) 0)
#endif

__device__ __forceinline__ V3 xyz(float4 v) { return v3(v.x, v.y, v.z); }

// One 64-byte geometry record (DBvhNode, DTri) into four float4. QB_LD256: two 256-bit loads (LDG.E.256, sm_100) instead of
// four 128-bit ones — every lane reads a different record, so each load instruction costs one L1 tag look-up per lane and the
// wide form halves them. QB_GEOM_EVICT_LAST: the records additionally stay in L2 ahead of the streaming wave state.
#ifndef QB_LD256
#define QB_LD256 1
#endif
#ifndef QB_GEOM_EVICT_LAST
#define QB_GEOM_EVICT_LAST 1
#endif
__device__ __forceinline__ void ld_rec64(const void *p, float4 &a, float4 &b, float4 &c, float4 &d)
{
#if QB_LD256
#if QB_GEOM_EVICT_LAST
#define QB_LD256_OP "ld.global.nc.L2::evict_last.v8.f32"
#else
#define QB_LD256_OP "ld.global.nc.v8.f32"
#endif
  asm(QB_LD256_OP " {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
      : "l"(p));
  asm(QB_LD256_OP " {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(c.x), "=f"(c.y), "=f"(c.z), "=f"(c.w), "=f"(d.x), "=f"(d.y), "=f"(d.z), "=f"(d.w)
      : "l"((const char *) p + 32));
#else
  const float4 *q = (const float4 *) p;
  a = __ldg(q);
  b = __ldg(q + 1);
  c = __ldg(q + 2);
  d = __ldg(q + 3);
#endif
}

// Node::ToNodeCoords(Ray): p' = itm*(p-pos); dir' = itm*((p+dir)-pos) - p'   (not itm*dir: the rounding differs)
__device__ __forceinline__ Ray3 to_node(const DNode *__restrict__ n, Ray3 r)
{
  const float4 i0 = __ldg(&n->i0), i1 = __ldg(&n->i1), i2 = __ldg(&n->i2);
  const V3 pos = v3(i0.w, i1.w, i2.w);
  const V3 a = r.p - pos;
  const V3 b = (r.p + r.d) - pos;
  Ray3 o;
  o.p = v3(i0.x * a.x + i1.x * a.y + i2.x * a.z, i0.y * a.x + i1.y * a.y + i2.y * a.z, i0.z * a.x + i1.z * a.y + i2.z * a.z);
  const V3 q = v3(i0.x * b.x + i1.x * b.y + i2.x * b.z, i0.y * b.x + i1.y * b.y + i2.y * b.z, i0.z * b.x + i1.z * b.y + i2.z * b.z);
  o.d = q - o.p;
  return o;
}

// Node::FromNodeCoords(HitInfo): p = tm*p + pos; N = normalize(itm^T * N)
__device__ __forceinline__ void from_node(const DNode *__restrict__ n, V3 &p, V3 &N)
{
  const float4 i0 = __ldg(&n->i0), i1 = __ldg(&n->i1), i2 = __ldg(&n->i2);
  const float4 t0 = __ldg(&n->t0), t1 = __ldg(&n->t1), t2 = __ldg(&n->t2);
  p = v3(t0.x * p.x + t1.x * p.y + t2.x * p.z, t0.y * p.x + t1.y * p.y + t2.y * p.z, t0.z * p.x + t1.z * p.y + t2.z * p.z) +
      v3(i0.w, i1.w, i2.w);
  N = normalize(v3(dot(xyz(i0), N), dot(xyz(i1), N), dot(xyz(i2), N)));
}

// ray of object k's node space: the root's transform was already applied to r0
__device__ __forceinline__ Ray3 chain_ray(const DScene &S, int k, Ray3 r0)
{
  const int b = __ldg(&S.chain_off[k]), e = __ldg(&S.chain_off[k + 1]);
  Ray3 r = r0;
  for (int c = b; c < e; ++c) r = to_node(&S.nodes[__ldg(&S.chain[c])], r);
  return r;
}

// ---- analytic objects: return the accepted t (z > t held) or a negative number ------------------------------
__device__ __forceinline__ float sphere_t(const Ray3 &r, float z)
{
  const float a = dot(r.d, r.d);
  const float b = 2.f * dot(r.p, r.d);
  const float c = dot(r.p, r.p) - 1;
  const float rcp2a = 1.f / (2.f * a);
  const float delta = b * b - 4 * a * c;
  float t = QB_BIGFLOAT;
  if (delta < 0) return -1.f;
  if (delta == 0) {
    const float t0 = -b * rcp2a;
    if (t0 <= QB_BIAS) return -1.f;
    t = t0;
  } else {
    const float sd = sqrtf(delta);
    const float t1 = (-b - sd) * rcp2a;
    const float t2 = (-b + sd) * rcp2a;
    if (t1 <= QB_BIAS && t2 <= QB_BIAS) return -1.f;
    else if (t1 > QB_BIAS) t = qmin(t, t1);
    else if (t2 > QB_BIAS) t = qmin(t, t2);
  }
  return z > t ? t : -1.f;
}

__device__ __forceinline__ float plane_t(const Ray3 &r, float z)
{
  const float dz = r.d.z;  // dot(dir,(0,0,1)) is exactly dir.z for finite dir
  if (qabs(dz) < 1e-7f) return -1.f;
  const float t = -r.p.z / dz;
  if (t <= QB_BIAS) return -1.f;
  if (!(z > t)) return -1.f;
  const V3 p = r.p + r.d * t;
  if (qabs(p.x) > 1.f || qabs(p.y) > 1.f) return -1.f;
  return t;
}

// Box::IntersectRay(ray, t_max) for the whole-mesh box (note the non-strict comparisons, unlike the BVH test).
// exit_out = where the ray leaves the box (QB_BIGFLOAT when every axis is degenerate).
__device__ __forceinline__ bool mesh_box_hit(const DMesh *__restrict__ m, const Ray3 &r, float t_max, float &exit_out, V3 &drcp)
{
  drcp = v3(1.f, 1.f, 1.f) / r.d;  // the same quotient TraceBVHNode forms (objects.cpp:333): returned for the traversal
  const V3 bmin = v3(__ldg(&m->bmin[0]), __ldg(&m->bmin[1]), __ldg(&m->bmin[2]));
  const V3 bmax = v3(__ldg(&m->bmax[0]), __ldg(&m->bmax[1]), __ldg(&m->bmax[2]));
  const V3 p0 = -(r.p - bmin) * drcp;
  const V3 p1 = -(r.p - bmax) * drcp;
  float t0x, t1x, t0y, t1y, t0z, t1z;
  if (qabs(r.d.x) < 1e-7f) { t0x = -QB_BIGFLOAT; t1x = QB_BIGFLOAT; } else { t0x = qmin(p0.x, p1.x); t1x = qmax(p0.x, p1.x); }
  if (qabs(r.d.y) < 1e-7f) { t0y = -QB_BIGFLOAT; t1y = QB_BIGFLOAT; } else { t0y = qmin(p0.y, p1.y); t1y = qmax(p0.y, p1.y); }
  if (qabs(r.d.z) < 1e-7f) { t0z = -QB_BIGFLOAT; t1z = QB_BIGFLOAT; } else { t0z = qmin(p0.z, p1.z); t1z = qmax(p0.z, p1.z); }
  const float entry = qmax(t0x, qmax(t0y, t0z));
  const float exit = qmin(t1x, qmin(t1y, t1z));
  exit_out = exit;
  return !(entry > t_max || entry > exit);
}

// IntersectTriangle's acceptance decision (objects.cpp:212-262). N is the unit face normal the reference recomputes per
// test, here computed once on the host with the same expression.
// The same decision without early exits: every lane of a converged LEAF phase runs one instruction stream. s and axis
// are the per-triangle constants 1 / TriangleArea(axis, A, B, C) and the dropped axis, computed on the host with the
// same expressions. Values computed past a failed check are never used (ok is false), so NaN/inf there are harmless.
__device__ __forceinline__ bool tri_accept(const Ray3 &r, float z, V3 A, V3 B, V3 C, V3 N, float s, int axis, float &t_out)
{
  const float dz = dot(r.d, N);
  const float pz = dot(r.p - A, N);
  const float t = -pz / dz;
  bool ok = !(qabs(dz) < 1e-7f) && !(t <= QB_BIAS) && (z > t);
  const V3 p = r.p + t * r.d;
  // coordinates kept by the projection: axis 0 -> (y, z), 1 -> (x, z), 2 -> (x, y)
  const bool a0 = axis == 0, a2 = axis == 2;
  const float pu = a0 ? p.y : p.x, pv = a2 ? p.y : p.z;
  const float Au = a0 ? A.y : A.x, Av = a2 ? A.y : A.z;
  const float Bu = a0 ? B.y : B.x, Bv = a2 ? B.y : B.z;
  const float Cu = a0 ? C.y : C.x, Cv = a2 ? C.y : C.z;
  const float a = ((Bu - pu) * (Cv - pv) - (Cu - pu) * (Bv - pv)) * s;   // TriangleArea(axis, p, B, C) * s
  const float b = ((Cu - pu) * (Av - pv) - (Au - pu) * (Cv - pv)) * s;   // TriangleArea(axis, p, C, A) * s
  const float c = 1.f - a - b;
  ok = ok && !(a < 0 || b < 0 || c < 0);
  t_out = t;
  return ok;
}

// IntersectTriangle without the comparison against the running hit distance (`hInfo.z > t`, objects.cpp:229): true when the
// triangle would be accepted by a ray whose hInfo.z is still beyond t. Lets a warp test the triangles of several leaves
// side by side; the owner of the ray applies `z > t` afterwards in the reference's order.
__device__ __forceinline__ bool tri_candidate(const Ray3 &r, V3 A, V3 B, V3 C, V3 N, float s, int axis, float &t_out)
{
  const float dz = dot(r.d, N);
  const float pz = dot(r.p - A, N);
  const float t = -pz / dz;
  bool ok = !(qabs(dz) < 1e-7f) && !(t <= QB_BIAS);
  const V3 p = r.p + t * r.d;
  const bool a0 = axis == 0, a2 = axis == 2;
  const float pu = a0 ? p.y : p.x, pv = a2 ? p.y : p.z;
  const float Au = a0 ? A.y : A.x, Av = a2 ? A.y : A.z;
  const float Bu = a0 ? B.y : B.x, Bv = a2 ? B.y : B.z;
  const float Cu = a0 ? C.y : C.x, Cv = a2 ? C.y : C.z;
  const float a = ((Bu - pu) * (Cv - pv) - (Cu - pu) * (Bv - pv)) * s;
  const float b = ((Cu - pu) * (Av - pv) - (Au - pu) * (Cv - pv)) * s;
  const float c = 1.f - a - b;
  ok = ok && !(a < 0 || b < 0 || c < 0);
  t_out = t;
  return ok;
}

// ---- hit attributes ---------------------------------------------------------------------------------------------
struct HitAttr {
  V3 p, N;          // world space after FromNodeCoords up the chain
  V3 uvw, duvw0, duvw1;
  int front, has_texture, mtl_id, face;  // face = id inside the mesh, -1 for sphere/plane
};

__device__ __forceinline__ V3 sphere_uvw(V3 p, float rcp_l)  // Sphere_TexCoord, objects.cpp:48-53
{
  const float RCP_PI = 1.f / 3.14159274101257324f, RCP_2PI = 1.f / (2.f * 3.14159274101257324f);
  return v3(0.5f - atan2f(p.x, p.y) * RCP_2PI, 0.5f + asinf(p.z * rcp_l) * RCP_PI, 0.f);
}
__device__ __forceinline__ V3 plane_uvw(V3 p) { return v3((p.x + 1.f) * 0.5f, (p.y + 1.f) * 0.5f, 0.f); }

// The texture part of an accepted hit on object node n (sphere, plane, or triangle WITH uv): what IntersectRay
// writes into hInfo.uvw / hInfo.duvw. rc/rx/ry are the rays in the node's space. Returns false for a triangle
// without texture vertices (the reference then leaves uvw/duvw/hasTexture as an earlier hit wrote them).
__device__ __forceinline__ bool hit_texture(const DScene &S, const DNode *__restrict__ n, int slot, float t, const Ray3 &rc,
                                            const Ray3 &rx, const Ray3 &ry, bool with_diff, V3 &uvw, V3 &d0, V3 &d1)
{
  const float rdx = 1.f / 0.01f;
  const int type = __ldg(&n->type);
  d0 = v3(0.f);
  d1 = v3(0.f);
  if (type == QG_OBJ_SPHERE) {
    const V3 p = rc.p + rc.d * t;
    const V3 N = normalize(p);
    uvw = sphere_uvw(p, 1.f);
    if (with_diff) {
      const float t_x = -dot(rx.p - p, N) / dot(rx.d, N);
      const float t_y = -dot(ry.p - p, N) / dot(ry.d, N);
      const V3 p_x = rx.p + rx.d * t_x, p_y = ry.p + ry.d * t_y;
      d0 = rdx * (sphere_uvw(p_x, 1.f / length(p_x)) - uvw);
      d1 = rdx * (sphere_uvw(p_y, 1.f / length(p_y)) - uvw);
    }
    return true;
  }
  if (type == QG_OBJ_PLANE) {
    const V3 p = rc.p + rc.d * t;
    uvw = plane_uvw(p);
    if (with_diff) {
      const float t_x = -rx.p.z / rx.d.z, t_y = -ry.p.z / ry.d.z;
      d0 = rdx * (plane_uvw(rx.p + rx.d * t_x) - uvw);
      d1 = rdx * (plane_uvw(ry.p + ry.d * t_y) - uvw);
    }
    return true;
  }
  const uint32_t gface = __ldg(&S.tri_face[slot]) & 0x7FFFFFFFu;
  const DFace *f = &S.faces[gface];
  if (!(__ldg(&f->mtl_uv) & 1)) return false;
  const DTri *tp = &S.tris[slot];
  const float4 ta = __ldg(&tp->a), tb = __ldg(&tp->b), tc = __ldg(&tp->c);
  const V3 A = xyz(ta), B = xyz(tb), C = xyz(tc), N = v3(ta.w, tb.w, tc.w);
  const int axis = tri_axis(N);
  const float s = 1.f / tri_area(axis, A, B, C);
  const V3 p = rc.p + t * rc.d;
  const float a = tri_area(axis, p, B, C) * s, b = tri_area(axis, p, C, A) * s, c = 1.f - a - b;
  const float u0 = __ldg(&f->uv[0]), v0 = __ldg(&f->uv[1]), u1 = __ldg(&f->uv[2]), v1 = __ldg(&f->uv[3]), u2 = __ldg(&f->uv[4]),
              v2 = __ldg(&f->uv[5]);
  uvw = v3(u0 * a + u1 * b + u2 * c, v0 * a + v1 * b + v2 * c, 0.f);  // GetTexCoord: VT0*bc.x + VT1*bc.y + VT2*bc.z
  if (with_diff) {
    const float t_x = -dot(rx.p - A, N) / dot(rx.d, N);
    const float t_y = -dot(ry.p - A, N) / dot(ry.d, N);
    const V3 p_x = rx.p + rx.d * t_x, p_y = ry.p + ry.d * t_y;
    const float ax = tri_area(axis, p_x, B, C) * s, bx = tri_area(axis, p_x, C, A) * s, cx = 1.f - ax - bx;
    const float ay = tri_area(axis, p_y, B, C) * s, by = tri_area(axis, p_y, C, A) * s, cy = 1.f - ay - by;
    d0 = rdx * (v3(u0 * ax + u1 * bx + u2 * cx, v0 * ax + v1 * bx + v2 * cx, 0.f) - uvw);
    d1 = rdx * (v3(u0 * ay + u1 * by + u2 * cy, v0 * ay + v1 * by + v2 * cy, 0.f) - uvw);
  }
  return true;
}

// Hit distances of the two differential rays on the accepted primitive's tangent plane (what IntersectRay stores in
// DiffHitInfo::x.z / y.z when ray.hasDiffRay): the Whitted-style material classes continue their differential rays from
// ray.x.p + ray.x.dir * hInfo.x.z (MtlBlinn_Basic.cpp:45-46). rc/rx/ry in the node's space.
__device__ __forceinline__ void diff_hit_t(const DScene &S, const DNode *__restrict__ n, int slot, float t, const Ray3 &rc, const Ray3 &rx,
                                           const Ray3 &ry, float &t_x, float &t_y)
{
  const int type = __ldg(&n->type);
  if (type == QG_OBJ_SPHERE) {
    const V3 p = rc.p + rc.d * t;
    const V3 N = normalize(p);
    t_x = -dot(rx.p - p, N) / dot(rx.d, N);
    t_y = -dot(ry.p - p, N) / dot(ry.d, N);
  } else if (type == QG_OBJ_PLANE) {
    t_x = -rx.p.z / rx.d.z;
    t_y = -ry.p.z / ry.d.z;
  } else {
    const DTri *tp = &S.tris[slot];
    const float4 ta = __ldg(&tp->a), tb = __ldg(&tp->b), tc = __ldg(&tp->c);
    const V3 A = xyz(ta), N = v3(ta.w, tb.w, tc.w);
    t_x = -dot(rx.p - A, N) / dot(rx.d, N);
    t_y = -dot(ry.p - A, N) / dot(ry.d, N);
  }
}

// Geometric part of the final hit (p, N, front, mtlID, face) in node space, then FromNodeCoords up to the root.
__device__ __forceinline__ void hit_geometry(const DScene &S, int k, const DNode *__restrict__ n, int slot, float t, const Ray3 &rc,
                                             HitAttr &h)
{
  const int type = __ldg(&n->type);
  V3 p, N;
  h.face = -1;
  h.mtl_id = 0;
  if (type == QG_OBJ_SPHERE) {
    p = rc.p + rc.d * t;
    N = normalize(p);
    h.front = dot(N, rc.d) <= 0;
  } else if (type == QG_OBJ_PLANE) {
    p = rc.p + rc.d * t;
    N = v3(0.f, 0.f, 1.f);
    h.front = rc.d.z <= 0;
  } else {
    const uint32_t gface = __ldg(&S.tri_face[slot]) & 0x7FFFFFFFu;
    const DFace *f = &S.faces[gface];
    const DTri *tp = &S.tris[slot];
    const float4 ta = __ldg(&tp->a), tb = __ldg(&tp->b), tc = __ldg(&tp->c);
    const V3 A = xyz(ta), B = xyz(tb), C = xyz(tc), Ng = v3(ta.w, tb.w, tc.w);
    h.front = dot(rc.d, Ng) <= 0;
    const int axis = tri_axis(Ng);
    const float s = 1.f / tri_area(axis, A, B, C);
    p = rc.p + t * rc.d;
    const float a = tri_area(axis, p, B, C) * s, b = tri_area(axis, p, C, A) * s, c = 1.f - a - b;
    // GetNormal: VN0*bc.x + VN1*bc.y + VN2*bc.z, left unnormalised until FromNodeCoords
    N = v3(__ldg(&f->n[0]) * a + __ldg(&f->n[3]) * b + __ldg(&f->n[6]) * c, __ldg(&f->n[1]) * a + __ldg(&f->n[4]) * b + __ldg(&f->n[7]) * c,
           __ldg(&f->n[2]) * a + __ldg(&f->n[5]) * b + __ldg(&f->n[8]) * c);
    h.mtl_id = (__ldg(&f->mtl_uv) >> 1) - 1;
    h.face = (int) (gface - __ldg(&S.meshes[__ldg(&n->mesh)].face_base));
  }
  // FromNodeCoords on the way out: the object node first, then its ancestors, then the root
  const int b = __ldg(&S.chain_off[k]), e = __ldg(&S.chain_off[k + 1]);
  for (int c = e - 1; c >= b; --c) from_node(&S.nodes[__ldg(&S.chain[c])], p, N);
  from_node(&S.nodes[0], p, N);
  h.p = p;
  h.N = N;
}

}  // namespace qb
