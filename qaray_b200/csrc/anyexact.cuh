// anyexact.cuh — the reference's OWN walk for any-hit rays the lane pool hands over (traverse.cuh, NEXT phase), one WARP per ray.
//
// Which rays: an any-hit ray without a far end (the shadow ray of a directional light, tmax = 1e30) whose direction has a
// degenerate axis (|d| < 1e-7) in a mesh's frame and which the culled walk found unoccluded. The reference's slab test ignores
// such an axis (src/objects/objects.cpp:353-376), so its walk visits every box the ray's other axes cross — tens of
// thousands of nodes — and can accept a rounding-noise hit at t ~ 1e6 on a triangle the ray never touches; to be exact all of
// those tests have to be made. Inside one lane that walk took up to 30 ms per ray while the rest of the warp idled: a 100x
// cliff for a directional light with a zero component over a mesh. Here the 32 lanes of a warp share the walk:
// a stack of node references in shared memory, every lane pops one, tests both children exactly as the reference does (no
// cull of any kind) or the triangles of a leaf, and pushes what is hit. For an any-hit ray the order does not matter: its
// distance bound never changes before the first accepted triangle, so "some reachable triangle accepts" is the same
// boolean whatever the order (TraceNodeShadow, src/scene/scene.cpp:35-46; GenLight::Shadow, src/lights/lights.cpp:39-48).
// Must be compiled with --fmad=false like trace.cuh.
#pragma once
#include "trace.cuh"

namespace qb {

#define QB_ANYX_STACK 2048  // node references per warp (8 KB); when it runs full the warp pops one node at a time (plain DFS)
#define QB_ANYX_WARPS 4

struct AnyQueue {
  uint32_t *q;      // ray indices
  uint32_t cap;
  uint32_t *count;  // pushed (may run past cap: clamp)
  uint32_t *fetch;  // work cursor of k_any_exact
};
__device__ __forceinline__ bool any_push(const AnyQueue &Q, uint32_t idx)
{
  if (Q.cap == 0u) return false;
  const uint32_t o = atomicAdd(Q.count, 1u);
  if (o >= Q.cap) return false;
  Q.q[o] = idx;
  return true;
}

// Box::IntersectRay of TraceBVHNode for one child box, degenerate axes ignored as the reference ignores them
__device__ __forceinline__ bool ref_box_hit(const Ray3 &r, const V3 &drcp, bool gx, bool gy, bool gz, float t_max, float bx0, float by0, float bz0,
                                            float bx1, float by1, float bz1)
{
  const float p0x = -(r.p.x - bx0) * drcp.x, p0y = -(r.p.y - by0) * drcp.y, p0z = -(r.p.z - bz0) * drcp.z;
  const float p1x = -(r.p.x - bx1) * drcp.x, p1y = -(r.p.y - by1) * drcp.y, p1z = -(r.p.z - bz1) * drcp.z;
  const float t0x = gx ? -QB_BIGFLOAT : qmin(p0x, p1x), t1x = gx ? QB_BIGFLOAT : qmax(p0x, p1x);
  const float t0y = gy ? -QB_BIGFLOAT : qmin(p0y, p1y), t1y = gy ? QB_BIGFLOAT : qmax(p0y, p1y);
  const float t0z = gz ? -QB_BIGFLOAT : qmin(p0z, p1z), t1z = gz ? QB_BIGFLOAT : qmax(p0z, p1z);
  const float entry = qmax(t0x, qmax(t0y, t0z));
  const float exit = qmin(t1x, qmin(t1y, t1z));
  return entry < t_max && entry < exit;
}

// one warp, one ray: true = occluded. stk = this warp's QB_ANYX_STACK words of shared memory.
__device__ __forceinline__ bool any_exact_ray(const DScene &S, const Ray3 &c, float tmax, uint32_t *stk, TraceCounters &cnt)
{
  const uint32_t FULL = 0xFFFFFFFFu;
  const int lane = threadIdx.x & 31;
  const Ray3 c0 = to_node(&S.nodes[0], c);
  for (int k = 0; k < S.num_obj; ++k) {  // every lane computes the object's ray: the same values everywhere
    const DNode *nd = &S.nodes[__ldg(&S.obj_node[k])];
    const Ray3 r = chain_ray(S, k, c0);
    const int type = __ldg(&nd->type);
    if (lane == 0) QB_CNT(objects);
    if (type != QG_OBJ_MESH) {
      const float t = type == QG_OBJ_SPHERE ? sphere_t(r, tmax) : plane_t(r, tmax);
      if (t > 0.f) return true;
      continue;
    }
    const DMesh *m = &S.meshes[__ldg(&nd->mesh)];
    if (__ldg(&m->num_faces) == 0u) continue;
    float mexit;
    V3 drcp;
    if (!mesh_box_hit(m, r, tmax, mexit, drcp)) continue;
    const bool gx = qabs(r.d.x) < 1e-7f, gy = qabs(r.d.y) < 1e-7f, gz = qabs(r.d.z) < 1e-7f;
    int sp = 1;  // warp-uniform
    if (lane == 0) stk[0] = __ldg(&m->root_ref);
    __syncwarp();
    while (sp > 0) {
      const int take = sp + 64 > QB_ANYX_STACK ? 1 : min(sp, 32);
      const bool mine = lane < take;
      const uint32_t ref = mine ? stk[sp - 1 - lane] : 0u;
      sp -= take;
      __syncwarp();
      uint32_t out0 = 0u, out1 = 0u;
      int n_out = 0;
      bool hit = false;
      if (mine) {
        if (ref & QB_LEAF_BIT) {
          const uint32_t base = ref & 0x0FFFFFFFu, count = ((ref >> 28) & 7u) + 1u;
          for (uint32_t i = 0; i < count && !hit; ++i) {
            const DTri *tp = &S.tris[base + i];
            const float4 ta = __ldg(&tp->a), tb = __ldg(&tp->b), tc = __ldg(&tp->c), td = __ldg(&tp->d);
            QB_CNT(tris);
            float t;
            hit = tri_accept(r, tmax, xyz(ta), xyz(tb), xyz(tc), v3(ta.w, tb.w, tc.w), td.x, __float_as_int(td.y), t);
          }
        } else {
          const DBvhNode *np = &S.bvh[ref];
          const float4 na = __ldg(&np->a), nb = __ldg(&np->b), nc = __ldg(&np->c);
          const uint4 ndr = __ldg(&np->d);
          QB_CNT(nodes);
          if (ref_box_hit(r, drcp, gx, gy, gz, tmax, na.x, na.y, na.z, na.w, nb.x, nb.y)) { out0 = ndr.x; n_out = 1; }
          if (ref_box_hit(r, drcp, gx, gy, gz, tmax, nb.z, nb.w, nc.x, nc.y, nc.z, nc.w)) {
            if (n_out) out1 = ndr.y; else out0 = ndr.y;
            ++n_out;
          }
        }
      }
      if (__any_sync(FULL, hit)) return true;
      int incl = n_out;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += v;
      }
      const int total = __shfl_sync(FULL, incl, 31);
      const int at = sp + incl - n_out;
      if (n_out > 0) stk[at] = out0;
      if (n_out > 1) stk[at + 1] = out1;
      sp += total;
      __syncwarp();
    }
  }
  return false;
}

// IO: load(idx, c, tmax), store_any(idx, occluded)
template <class IO> __device__ __forceinline__ void any_exact_pass(const DScene &S, IO &io, const AnyQueue &Q, uint32_t *s_stack, TraceCounters &cnt)
{
  const uint32_t n = min(*Q.count, Q.cap);
  const int lane = threadIdx.x & 31;
  uint32_t *stk = s_stack + (threadIdx.x >> 5) * QB_ANYX_STACK;
  for (;;) {
    uint32_t o = 0;
    if (lane == 0) o = atomicAdd(Q.fetch, 1u);
    o = __shfl_sync(0xFFFFFFFFu, o, 0);
    if (o >= n) break;
    const uint32_t idx = Q.q[o];
    Ray3 c;
    float tmax;
    if (!io.load(idx, c, tmax)) continue;  // every lane loads the same ray
    const bool occluded = any_exact_ray(S, c, tmax, stk, cnt);
    if (lane == 0) io.store_any(idx, occluded);
    __syncwarp();
  }
}

}  // namespace qb
