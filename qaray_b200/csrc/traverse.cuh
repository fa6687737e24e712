// traverse.cuh — the traversal core of every trace kernel: a warp is a POOL of 32 lanes, each lane walks its own ray
// through Scene::TraceNodeNormal / TraceNodeShadow (src/scene/scene.cpp:35-74) as one flat state machine
//
//     object k of the pre-order object list -> (Node::ToNodeCoords chain) -> Sphere | Plane | TriObj box + BVH
//
// instead of nested loops, so lanes of one warp can be in different objects / BVH depths at the same time, and a lane
// whose ray is finished takes the next ray from the device cursor while its neighbours keep going (persistent
// threads with dynamic fetch). The per-ray ORDER of node visits, triangle tests and comparisons is exactly the
// reference's (TriObj::TraceBVHNode, src/objects/objects.cpp:324-420: near child first, far child deferred), so
// hit/miss, primitive id and t stay bit-identical; only the SIMD scheduling differs.
//
// BVH stack: the near child stays in a register, only deferred far children are stored, in SHARED memory laid out
// [entry][thread] (conflict-free). Depth is bounded by the tree depth, checked on the host at upload (QB_STACK).
// Must be compiled with --fmad=false like trace.cuh.
#pragma once
#include "trace.cuh"

namespace qb {

#define QB_REF_EMPTY 0x7FFFFFFFu   // "no current node" (an internal index never reaches 2^31-1)
#ifndef QB_INNER_REPS
#define QB_INNER_REPS 3            // node visits per vote while the inner group holds together
#endif

// what the closest-hit traversal leaves per ray; attributes are computed later by finish_hit (full warps)
struct HitRec {
  float z;
  int k, slot;          // object index (-1 miss), triangle slot (-1 sphere/plane)
  int tex_k, tex_slot;  // last hit that wrote texture coordinates (stale carry-over), its distance
  float tex_t;
  int last_tri;         // last accepted triangle of any mesh (stale mtlID)
};

// IO concept:
//   uint32_t count() const                       number of rays
//   uint32_t *cursor()                           device work cursor
//   bool load(uint32_t idx, Ray3 &c, float &tmax)    world-space ray; false = padding entry (skipped, store not called)
//   void store(uint32_t idx, const HitRec &)     closest
//   void store_any(uint32_t idx, bool occluded)  any-hit
template <bool ANY, class IO>
__device__ __forceinline__ void traverse_pool(const DScene &S, IO &io, uint32_t *__restrict__ s_stack, TraceCounters &cnt)
{
  const int lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t n = io.count();
  const DNode *root = &S.nodes[0];
  uint32_t *stk = s_stack + threadIdx.x;  // entry e at stk[e * QB_BLOCK]

  bool active = false, exhausted = false;
  uint32_t idx = 0;
  Ray3 c0, r;
  V3 drcp = v3(0.f);
  bool gx = false, gy = false, gz = false;
  float mesh_exit = 0.f;  // where the ray leaves the current mesh's box (bounds the degenerate-axis cull below)
  float z_entry = 0.f;    // h.z when the current mesh was entered
  float z_sane = 0.f;     // hits of the current mesh beyond this distance can only be rounding artefacts (see the cull below)
  int cull_state = 0;     // 0 nothing dropped by the cull yet, 1 something dropped, 2 cull off (exact repeat of the mesh)
  int k = 0, sp = 0;
  uint32_t cur = QB_REF_EMPTY;
  HitRec h;
  h.z = 0.f; h.k = -1; h.slot = -1; h.tex_k = -1; h.tex_slot = -1; h.tex_t = 0.f; h.last_tri = -1;
  int uv_slot = -1;   // last accepted triangle with uv inside the CURRENT object
  float uv_t = 0.f;
  int obj_slot = -1;  // last accepted triangle inside the current object

  // Every lane is in one of four states; each iteration the warp votes and runs ONE phase, the one most lanes wait
  // for, with all of those lanes converged in it. Lanes in other states keep their state until their phase wins.
  //   FETCH  no ray: take the next one from the cursor            INNER  current node is an internal BVH node
  //   LEAF   current node is a leaf (<= 4 triangles)              NEXT   object finished: report it, set up the next
  for (;;) {
    const bool s_fetch = !active && !exhausted;
    const bool s_inner = active && cur != QB_REF_EMPTY && !(cur & QB_LEAF_BIT);
    const bool s_leaf = active && (cur & QB_LEAF_BIT) != 0u;
    const bool s_next = active && cur == QB_REF_EMPTY;
    const uint32_t m_fetch = __ballot_sync(0xFFFFFFFFu, s_fetch), m_inner = __ballot_sync(0xFFFFFFFFu, s_inner);
    const uint32_t m_leaf = __ballot_sync(0xFFFFFFFFu, s_leaf), m_next = __ballot_sync(0xFFFFFFFFu, s_next);
    if ((m_fetch | m_inner | m_leaf | m_next) == 0u) break;
    const int n_fetch = __popc(m_fetch), n_inner = __popc(m_inner), n_leaf = __popc(m_leaf), n_next = __popc(m_next);

    if (n_inner >= n_leaf && n_inner >= n_next && n_inner >= n_fetch) {
      // ---- INNER: node visits (both children's boxes); the near child becomes current, the far child is deferred.
      // Up to QB_INNER_REPS visits per vote while at least half of the lanes that started are still on inner nodes.
#pragma unroll 1
      for (int rep = 0; rep < QB_INNER_REPS; ++rep) {
        const bool go = active && cur != QB_REF_EMPTY && !(cur & QB_LEAF_BIT);
        if (rep > 0 && 2 * __popc(__ballot_sync(0xFFFFFFFFu, go)) < n_inner) break;
        if (go) {
          const DBvhNode *np = &S.bvh[cur];
          const float4 na = __ldg(&np->a), nb = __ldg(&np->b), nc = __ldg(&np->c);
          const uint4 ndr = __ldg(&np->d);
          QB_CNT(nodes);
          const float p00x = -(r.p.x - na.x) * drcp.x, p00y = -(r.p.y - na.y) * drcp.y, p00z = -(r.p.z - na.z) * drcp.z;
          const float p01x = -(r.p.x - na.w) * drcp.x, p01y = -(r.p.y - nb.x) * drcp.y, p01z = -(r.p.z - nb.y) * drcp.z;
          const float p10x = -(r.p.x - nb.z) * drcp.x, p10y = -(r.p.y - nb.w) * drcp.y, p10z = -(r.p.z - nc.x) * drcp.z;
          const float p11x = -(r.p.x - nc.y) * drcp.x, p11y = -(r.p.y - nc.z) * drcp.y, p11z = -(r.p.z - nc.w) * drcp.z;
          const float t_max = h.z;
          float entry0, entry1, exit0, exit1;
          bool h0, h1;
          if (!(gx || gy || gz)) {
            // no degenerate axis: every p is finite or +-inf, never NaN, so the reference's MIN/MAX macros
            // (x < y ? x : y) and fminf/fmaxf pick the same value up to the sign of a zero, which no comparison sees
            entry0 = fmaxf(fminf(p00x, p01x), fmaxf(fminf(p00y, p01y), fminf(p00z, p01z)));
            entry1 = fmaxf(fminf(p10x, p11x), fmaxf(fminf(p10y, p11y), fminf(p10z, p11z)));
            exit0 = fminf(fmaxf(p00x, p01x), fminf(fmaxf(p00y, p01y), fmaxf(p00z, p01z)));
            exit1 = fminf(fmaxf(p10x, p11x), fminf(fmaxf(p10y, p11y), fmaxf(p10z, p11z)));
            h0 = (entry0 < t_max && entry0 < exit0);
            h1 = (entry1 < t_max && entry1 < exit1);
          } else {
            const float t00x = gx ? -QB_BIGFLOAT : qmin(p00x, p01x), t01x = gx ? QB_BIGFLOAT : qmax(p00x, p01x);
            const float t10x = gx ? -QB_BIGFLOAT : qmin(p10x, p11x), t11x = gx ? QB_BIGFLOAT : qmax(p10x, p11x);
            const float t00y = gy ? -QB_BIGFLOAT : qmin(p00y, p01y), t01y = gy ? QB_BIGFLOAT : qmax(p00y, p01y);
            const float t10y = gy ? -QB_BIGFLOAT : qmin(p10y, p11y), t11y = gy ? QB_BIGFLOAT : qmax(p10y, p11y);
            const float t00z = gz ? -QB_BIGFLOAT : qmin(p00z, p01z), t01z = gz ? QB_BIGFLOAT : qmax(p00z, p01z);
            const float t10z = gz ? -QB_BIGFLOAT : qmin(p10z, p11z), t11z = gz ? QB_BIGFLOAT : qmax(p10z, p11z);
            entry0 = qmax(t00x, qmax(t00y, t00z)); entry1 = qmax(t10x, qmax(t10y, t10z));
            exit0 = qmin(t01x, qmin(t01y, t01z)); exit1 = qmin(t11x, qmin(t11y, t11z));
            h0 = (entry0 < t_max && entry0 < exit0);
            h1 = (entry1 < t_max && entry1 < exit1);
            // Degenerate axis (|d| < 1e-7): the reference's slab test ignores that axis altogether, so such a ray would
            // walk every box its other two axes cross (the exact image-centre ray of sample 0 visits the whole tree, one
            // dependent load after another: a 20 ms tail). A box the ray cannot reach along that axis holds no
            // acceptable triangle: an accepted hit point lies inside its triangle's bounds up to rounding, and the ray's
            // coordinate moves by at most |d| * t with t < min(z, exit of the mesh box). Dropping such boxes changes no
            // hit, no distance and no order among the boxes that remain. Margin: twice that travel + 1e-4 relative.
            // One exception is handled when the mesh is finished (NEXT phase): a triangle almost parallel to the ray gives
            // t = -pz/dz in the millions, where the barycentric test degenerates to rounding noise and can ACCEPT a point
            // far outside the triangle (seen: t = 1.9e6 in a 50-unit scene, tests/golden/edge_trc_scene_tower.npz). Such a
            // hit lies far beyond the mesh box, and it can only surface in the reference's result when nothing nearer was
            // hit. So if something was dropped and the ray leaves the mesh with z still beyond z_sane = max(box exit,
            // 1000 box diagonals) — in practice: still without any hit — the mesh is traversed again with the cull off
            // (cull_state 2), exactly as the reference walks it. A ray that already holds a real hit never pays this.
            const float t_far = qmin(t_max, mesh_exit);
            const bool h0r = h0, h1r = h1;
            if (cull_state != 2) {
            if (gx) {
              const float mg = 2.f * qabs(r.d.x) * t_far;
              const float m0 = mg + 1e-4f * qmax(qabs(r.p.x), qmax(qabs(na.x), qabs(na.w)));
              const float m1 = mg + 1e-4f * qmax(qabs(r.p.x), qmax(qabs(nb.z), qabs(nc.y)));
              h0 = h0 && !(r.p.x < na.x - m0 || r.p.x > na.w + m0);
              h1 = h1 && !(r.p.x < nb.z - m1 || r.p.x > nc.y + m1);
            }
            if (gy) {
              const float mg = 2.f * qabs(r.d.y) * t_far;
              const float m0 = mg + 1e-4f * qmax(qabs(r.p.y), qmax(qabs(na.y), qabs(nb.x)));
              const float m1 = mg + 1e-4f * qmax(qabs(r.p.y), qmax(qabs(nb.w), qabs(nc.z)));
              h0 = h0 && !(r.p.y < na.y - m0 || r.p.y > nb.x + m0);
              h1 = h1 && !(r.p.y < nb.w - m1 || r.p.y > nc.z + m1);
            }
            if (gz) {
              const float mg = 2.f * qabs(r.d.z) * t_far;
              const float m0 = mg + 1e-4f * qmax(qabs(r.p.z), qmax(qabs(na.z), qabs(nb.y)));
              const float m1 = mg + 1e-4f * qmax(qabs(r.p.z), qmax(qabs(nc.x), qabs(nc.w)));
              h0 = h0 && !(r.p.z < na.z - m0 || r.p.z > nb.y + m0);
              h1 = h1 && !(r.p.z < nc.x - m1 || r.p.z > nc.w + m1);
            }
            if ((h0r && !h0) || (h1r && !h1)) cull_state = 1;  // something the reference would visit was dropped
            }
          }
          {
            // the three outcomes (both children, one, none) as selects + one predicated store / load, so the lanes of the
            // phase stay converged instead of running three serial branches
            const bool both = h0 && h1, none = !(h0 || h1);
            const bool first0 = entry0 < entry1;  // the reference pushes the other child first, then this one
            const bool take0 = both ? first0 : h0;
            const uint32_t near_ref = take0 ? ndr.x : ndr.y, far_ref = take0 ? ndr.y : ndr.x;
            if (both) stk[sp * QB_BLOCK] = far_ref;
            sp += both ? 1 : 0;
            const bool pop = none && sp > 0;
            sp -= pop ? 1 : 0;
            uint32_t popped = QB_REF_EMPTY;
            if (pop) popped = stk[sp * QB_BLOCK];
            cur = none ? popped : near_ref;
          }
        }
      }
    } else if (n_leaf >= n_next && n_leaf >= n_fetch) {
      // ---- LEAF: the leaf's triangles in elements[] order, branch-free per triangle ------------------------------------
      if (s_leaf) {
        const uint32_t base = cur & 0x0FFFFFFFu;
        const uint32_t count = ((cur >> 28) & 7u) + 1u;
        cur = sp > 0 ? stk[(--sp) * QB_BLOCK] : QB_REF_EMPTY;
        for (uint32_t i = 0; i < count; ++i) {
          const DTri *tp = &S.tris[base + i];
          const float4 ta = __ldg(&tp->a), tb = __ldg(&tp->b), tc = __ldg(&tp->c), td = __ldg(&tp->d);
          QB_CNT(tris);
          float t;
          if (tri_accept(r, h.z, xyz(ta), xyz(tb), xyz(tc), v3(ta.w, tb.w, tc.w), td.x, __float_as_int(td.y), t)) {
            if (ANY) { io.store_any(idx, true); active = false; break; }
            h.z = t;
            obj_slot = (int) (base + i);
            if (__float_as_uint(td.z) & 0x80000000u) { uv_slot = obj_slot; uv_t = t; }
          }
        }
      }
    } else if (n_next >= n_fetch) {
      // ---- NEXT: the object just traversed reports its result, then the next object of the list is set up ----------
      bool redo = false;
      if (s_next && k > 0 && cull_state == 1) {
        // the cull dropped boxes and the mesh produced no hit inside its own box: repeat it exactly as the reference walks it
        cull_state = 0;
        // closest-hit rays of the wavefront hand the check to k_redo_detect / k_redo_exact (trace_kernels.cu) and go on: a
        // single lane repeating a whole mesh costs up to ~30 ms, and some sample ranges hold a few dozen such rays
        if (!(h.z <= z_sane) && (ANY || !io.defer(idx, k - 1, z_entry))) {
          redo = true;
          cull_state = 2;
          h.z = z_entry;
          obj_slot = -1; uv_slot = -1; sp = 0;
          cur = __ldg(&S.meshes[__ldg(&S.nodes[__ldg(&S.obj_node[k - 1])].mesh)].root_ref);
        }
      }
      if (s_next && !redo) {
        if (k > 0 && obj_slot >= 0) {
          h.k = k - 1; h.slot = obj_slot; h.last_tri = obj_slot;
          if (uv_slot >= 0) { h.tex_k = k - 1; h.tex_slot = uv_slot; h.tex_t = uv_t; }
          obj_slot = -1;
        }
        if (k == S.num_obj) {
          if (ANY) io.store_any(idx, false); else io.store(idx, h);
          active = false;
        } else {
          const DNode *nd = &S.nodes[__ldg(&S.obj_node[k])];
          r = chain_ray(S, k, c0);
          const int type = __ldg(&nd->type);
          QB_CNT(objects);
          ++k;
          obj_slot = -1; uv_slot = -1;
          if (type == QG_OBJ_MESH) {
            const DMesh *m = &S.meshes[__ldg(&nd->mesh)];
            if (__ldg(&m->num_faces) != 0u && mesh_box_hit(m, r, h.z, mesh_exit, drcp)) {
              cur = __ldg(&m->root_ref);
              cull_state = io.exact(idx, k - 1) ? 2 : 0;  // k was advanced above: k - 1 is this object
              z_entry = h.z;
              {
                const V3 ext = v3(__ldg(&m->bmax[0]) - __ldg(&m->bmin[0]), __ldg(&m->bmax[1]) - __ldg(&m->bmin[1]),
                                  __ldg(&m->bmax[2]) - __ldg(&m->bmin[2]));
                z_sane = qmax(mesh_exit * 1.001f, 1000.f * length(ext));
              }
              gx = qabs(r.d.x) < 1e-7f; gy = qabs(r.d.y) < 1e-7f; gz = qabs(r.d.z) < 1e-7f;
            }
          } else {
            const float t = type == QG_OBJ_SPHERE ? sphere_t(r, h.z) : plane_t(r, h.z);
            if (t > 0.f) {
              if (ANY) {
                io.store_any(idx, true);
                active = false;
              } else {
                h.z = t; h.k = k - 1; h.slot = -1;
                h.tex_k = k - 1; h.tex_slot = -1; h.tex_t = t;  // sphere / plane always write uvw
              }
            }
          }
        }
      }
    } else {
      // ---- FETCH: idle lanes take consecutive rays from the device cursor (one atomic per warp) -------------------------
      uint32_t base = 0;
      if (lane == 0) base = atomicAdd(io.cursor(), (uint32_t) n_fetch);
      base = __shfl_sync(0xFFFFFFFFu, base, 0);
      if (base + (uint32_t) n_fetch >= n) exhausted = true;
      if (s_fetch) {
        const uint32_t my = base + (uint32_t) __popc(m_fetch & lt_mask);
        float tmax = QB_BIGFLOAT;
        Ray3 c;
        const uint32_t ray = my < n ? io.map(my) : my;
        if (my < n && io.load(ray, c, tmax)) {
          idx = ray;
          c0 = to_node(root, c);
          h.z = tmax; h.k = -1; h.slot = -1; h.tex_k = -1; h.tex_slot = -1; h.tex_t = 0.f; h.last_tri = -1;
          k = 0; sp = 0; cur = QB_REF_EMPTY; obj_slot = -1; uv_slot = -1;
          active = true;
        }
      }
    }
  }
}

// Attributes of a finished closest-hit ray (what Scene::TraceNodeNormal leaves in DiffHitInfo::c): the tail of
// trace_closest, run by full warps after the traversal. c/x/y world-space rays as in trace_closest.
__device__ __forceinline__ void finish_hit(const DScene &S, const HitRec &rec, const Ray3 &c, const Ray3 &x, const Ray3 &y, bool with_diff,
                                           bool stale_tex, HitAttr &h, float *txy = nullptr)
{
  const DNode *root = &S.nodes[0];
  const Ray3 c0 = to_node(root, c);
  const int best_k = rec.k, best_slot = rec.slot;
  const DNode *n = &S.nodes[__ldg(&S.obj_node[best_k])];
  const Ray3 r = chain_ray(S, best_k, c0);
  hit_geometry(S, best_k, n, best_slot, rec.z, r, h);
  if (txy) {  // DiffHitInfo::x.z / y.z of the accepted hit (objects.cpp:113-131,180-196,264-296)
    txy[0] = txy[1] = rec.z;
    if (with_diff) diff_hit_t(S, n, best_slot, rec.z, r, chain_ray(S, best_k, to_node(root, x)), chain_ray(S, best_k, to_node(root, y)), txy[0], txy[1]);
  }
  if (best_slot < 0 && rec.last_tri >= 0 && stale_tex)
    h.mtl_id = (__ldg(&S.faces[__ldg(&S.tri_face[rec.last_tri]) & 0x7FFFFFFFu].mtl_uv) >> 1) - 1;
  h.uvw = v3(0.5f, 0.5f, 0.5f);
  h.duvw0 = v3(0.f);
  h.duvw1 = v3(0.f);
  h.has_texture = 0;
  if (rec.tex_k == best_k && rec.tex_slot == best_slot) {
    Ray3 rx = r, ry = r;
    if (with_diff) {
      rx = chain_ray(S, best_k, to_node(root, x));
      ry = chain_ray(S, best_k, to_node(root, y));
    }
    h.has_texture = hit_texture(S, n, best_slot, rec.z, r, rx, ry, with_diff, h.uvw, h.duvw0, h.duvw1) ? 1 : 0;
  } else if (rec.tex_k >= 0 && stale_tex) {
    const DNode *tn = &S.nodes[__ldg(&S.obj_node[rec.tex_k])];
    const Ray3 tr = chain_ray(S, rec.tex_k, c0);
    Ray3 rx = tr, ry = tr;
    if (with_diff) {
      rx = chain_ray(S, rec.tex_k, to_node(root, x));
      ry = chain_ray(S, rec.tex_k, to_node(root, y));
    }
    h.has_texture = hit_texture(S, tn, rec.tex_slot, rec.tex_t, tr, rx, ry, with_diff, h.uvw, h.duvw0, h.duvw1) ? 1 : 0;
  }
}

}  // namespace qb
