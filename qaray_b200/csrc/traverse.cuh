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

#ifndef QB_STACK_ASM
#define QB_STACK_ASM 1
#endif
// shared memory through 32-bit shared-space addresses (see QB_STACK_ASM in traverse_pool); volatile keeps the accesses of one
// thread in program order, which is all the thread-private words need
__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v)); }
__device__ __forceinline__ uint32_t lds_u32(uint32_t a)
{
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ float4 lds_f4(uint32_t a)
{
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}

#ifndef QB_COOP_LEAF
#define QB_COOP_LEAF 1
#endif
#ifndef QB_W_LEAF
#define QB_W_LEAF 8
#endif
// vote weights (INNER counts 4 per lane). Re-tuned on the final kernels, ms per step tower / xmas: NEXT 6 + FETCH 4 (before)
// 71.8 / 419.4; FETCH 2: 71.2 / 404.6; NEXT 4: 71.5 / 414.6; FETCH 2 + NEXT 5: 71.3 / 402.8; + NEXT 4: 71.85 / 401.0; + NEXT 3:
// 71.7 / 398.9; FETCH 1: 72.7; LEAF 6 / 10 / 12: 72.7 / 71.75 / 71.8
#ifndef QB_W_NEXT
#define QB_W_NEXT 5
#endif
#ifndef QB_W_FETCH
#define QB_W_FETCH 2
#endif
#ifndef QB_W_LEAF_ANY  // the any-hit kernels' own weights
#define QB_W_LEAF_ANY QB_W_LEAF
#endif
#ifndef QB_W_NEXT_ANY
#define QB_W_NEXT_ANY QB_W_NEXT
#endif
#ifndef QB_W_FETCH_ANY
#define QB_W_FETCH_ANY QB_W_FETCH
#endif
#define QB_COOP_SLOTS 128          // (ray, triangle) pairs a warp's LEAF phase handles per round
#define QB_REF_EMPTY 0x7FFFFFFFu   // "no current node" (an internal index never reaches 2^31-1)
#ifndef QB_NEXT_REPS
#define QB_NEXT_REPS 4             // objects per vote while the group between objects holds together
#endif
#ifndef QB_INNER_REPS
#define QB_INNER_REPS 5            // node visits per vote while the inner group holds together (A/B: 3 -> 5 -0.9 ms per tower step)
#endif

// what the closest-hit traversal leaves per ray; attributes are computed later by finish_hit (full warps)
struct HitRec {
  float z;
  int k, slot;          // object index (-1 miss), triangle slot (-1 sphere/plane)
  int tex_k, tex_slot;  // last hit that wrote texture coordinates (stale carry-over), its distance
  float tex_t;
  int last_tri;         // last accepted triangle of any mesh (stale mtlID)
};

// ---- per-block scene table ------------------------------------------------------------------------------------------
// Setting up the next object of a ray (NEXT phase) used to chase five dependent global loads (obj_node -> nodes ->
// chain_off -> chain -> node matrix -> mesh record) at 12-14 of 32 lanes: a quarter of the kernel's instructions and
// most of its long-scoreboard stalls per ray. The object list is tiny (tower 6, xmas 15 objects), so every block keeps
// it in shared memory: the ToNodeCoords matrices of the first QB_TAB_NODES scene nodes, and one 48-byte record per object
// (mesh box, root ref, the cull's sanity distance, type, ancestor chain). Scenes larger than the table read the rest
// from global memory through the same accessors.
#ifndef QB_TAB_NODES
#define QB_TAB_NODES 24
#endif
#ifndef QB_TAB_OBJS
#define QB_TAB_OBJS 24
#endif
#define QB_TAB_CHAIN 64
#define QB_NO_FACES 0x7FFFFFFEu  // root_ref of a mesh without triangles (TriObj::IntersectRay never reaches the BVH)

struct ObjRec {
  float4 lo;  // mesh bmin.xyz, root_ref (uint bits)
  float4 hi;  // mesh bmax.xyz, 1000 * length(bmax - bmin)
  int4 id;    // type, -, chain begin, chain end (into the chain list: ancestors root-first, root excluded, object node last)
};

struct SceneTab {
  float4 node[QB_TAB_NODES][3];  // i0, i1, i2 of DNode
  ObjRec obj[QB_TAB_OBJS];
  int chain[QB_TAB_CHAIN];
  int chain_ok;                  // every chain entry of the first QB_TAB_OBJS objects is in chain[]
};

__device__ __forceinline__ ObjRec make_obj_rec(const DScene &S, int k)
{
  ObjRec o;
  const DNode *nd = &S.nodes[__ldg(&S.obj_node[k])];
  const int type = __ldg(&nd->type);
  o.id = make_int4(type, 0, __ldg(&S.chain_off[k]), __ldg(&S.chain_off[k + 1]));
  o.lo = make_float4(0.f, 0.f, 0.f, __uint_as_float(QB_NO_FACES));
  o.hi = make_float4(0.f, 0.f, 0.f, 0.f);
  if (type == QG_OBJ_MESH) {
    const DMesh *m = &S.meshes[__ldg(&nd->mesh)];
    const V3 bmin = v3(__ldg(&m->bmin[0]), __ldg(&m->bmin[1]), __ldg(&m->bmin[2]));
    const V3 bmax = v3(__ldg(&m->bmax[0]), __ldg(&m->bmax[1]), __ldg(&m->bmax[2]));
    const uint32_t root = __ldg(&m->num_faces) != 0u ? __ldg(&m->root_ref) : QB_NO_FACES;
    o.lo = make_float4(bmin.x, bmin.y, bmin.z, __uint_as_float(root));
    o.hi = make_float4(bmax.x, bmax.y, bmax.z, 1000.f * length(bmax - bmin));
  }
  return o;
}

// all threads of the block; ends with __syncthreads()
__device__ __forceinline__ void load_scene_tab(const DScene &S, SceneTab &T)
{
  for (int i = threadIdx.x; i < min(S.num_nodes, QB_TAB_NODES); i += blockDim.x) {
    T.node[i][0] = __ldg(&S.nodes[i].i0);
    T.node[i][1] = __ldg(&S.nodes[i].i1);
    T.node[i][2] = __ldg(&S.nodes[i].i2);
  }
  const int no = min(S.num_obj, QB_TAB_OBJS);
  for (int k = threadIdx.x; k < no; k += blockDim.x) T.obj[k] = make_obj_rec(S, k);
  const int nc = __ldg(&S.chain_off[no]);
  for (int c = threadIdx.x; c < min(nc, QB_TAB_CHAIN); c += blockDim.x) T.chain[c] = __ldg(&S.chain[c]);
  if (threadIdx.x == 0) T.chain_ok = nc <= QB_TAB_CHAIN;
  __syncthreads();
}

__device__ __forceinline__ Ray3 to_node_q(float4 i0, float4 i1, float4 i2, Ray3 r)  // to_node on loaded matrix rows
{
  const V3 pos = v3(i0.w, i1.w, i2.w);
  const V3 a = r.p - pos;
  const V3 b = (r.p + r.d) - pos;
  Ray3 o;
  o.p = v3(i0.x * a.x + i1.x * a.y + i2.x * a.z, i0.y * a.x + i1.y * a.y + i2.y * a.z, i0.z * a.x + i1.z * a.y + i2.z * a.z);
  const V3 q = v3(i0.x * b.x + i1.x * b.y + i2.x * b.z, i0.y * b.x + i1.y * b.y + i2.y * b.z, i0.z * b.x + i1.z * b.y + i2.z * b.z);
  o.d = q - o.p;
  return o;
}
#if QB_STACK_ASM
// the block's table through its 32-bit shared address (TabRef = that address; see QB_STACK_ASM)
typedef uint32_t TabRef;
__device__ __forceinline__ TabRef tab_ref(const SceneTab &T) { return smem_addr(&T); }
__device__ __forceinline__ int tab_chain(TabRef T, int c) { return (int) lds_u32(T + (uint32_t) offsetof(SceneTab, chain) + 4u * (uint32_t) c); }
__device__ __forceinline__ bool tab_chain_ok(TabRef T) { return lds_u32(T + (uint32_t) offsetof(SceneTab, chain_ok)) != 0u; }
__device__ __forceinline__ Ray3 to_node_tab(const DScene &S, TabRef T, int n, Ray3 r)
{
  if (n < QB_TAB_NODES) {
    const uint32_t a = T + (uint32_t) offsetof(SceneTab, node) + 48u * (uint32_t) n;
    return to_node_q(lds_f4(a), lds_f4(a + 16u), lds_f4(a + 32u), r);
  }
  return to_node(&S.nodes[n], r);
}
__device__ __forceinline__ ObjRec obj_rec(const DScene &S, TabRef T, int k)
{
  if (k < QB_TAB_OBJS) {
    const uint32_t a = T + (uint32_t) offsetof(SceneTab, obj) + (uint32_t) sizeof(ObjRec) * (uint32_t) k;
    ObjRec o;
    o.lo = lds_f4(a);
    o.hi = lds_f4(a + 16u);
    const float4 id = lds_f4(a + 32u);
    o.id = make_int4(__float_as_int(id.x), __float_as_int(id.y), __float_as_int(id.z), __float_as_int(id.w));
    return o;
  }
  return make_obj_rec(S, k);
}
#else
typedef const SceneTab &TabRef;
__device__ __forceinline__ TabRef tab_ref(const SceneTab &T) { return T; }
__device__ __forceinline__ int tab_chain(TabRef T, int c) { return T.chain[c]; }
__device__ __forceinline__ bool tab_chain_ok(TabRef T) { return T.chain_ok != 0; }
__device__ __forceinline__ Ray3 to_node_tab(const DScene &S, TabRef T, int n, Ray3 r)
{
  if (n < QB_TAB_NODES) return to_node_q(T.node[n][0], T.node[n][1], T.node[n][2], r);
  return to_node(&S.nodes[n], r);
}
__device__ __forceinline__ ObjRec obj_rec(const DScene &S, TabRef T, int k)
{
  if (k < QB_TAB_OBJS) return T.obj[k];
  return make_obj_rec(S, k);
}
#endif
// ray of object k's node space (the root's transform already applied to r0): chain_ray through the table
__device__ __forceinline__ Ray3 chain_ray_tab(const DScene &S, TabRef T, const ObjRec &o, int k, Ray3 r0)
{
  Ray3 r = r0;
  if (k < QB_TAB_OBJS && tab_chain_ok(T)) {
    for (int c = o.id.z; c < o.id.w; ++c) r = to_node_tab(S, T, tab_chain(T, c), r);
  } else {
    for (int c = o.id.z; c < o.id.w; ++c) r = to_node_tab(S, T, __ldg(&S.chain[c]), r);
  }
  return r;
}
// Box::IntersectRay of mesh_box_hit on the table's box
__device__ __forceinline__ bool mesh_box_hit_q(const ObjRec &o, const Ray3 &r, float t_max, float &exit_out, V3 &drcp)
{
  drcp = v3(1.f, 1.f, 1.f) / r.d;
  const V3 bmin = v3(o.lo.x, o.lo.y, o.lo.z), bmax = v3(o.hi.x, o.hi.y, o.hi.z);
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

// Boxes behind the ray. The reference's slab test accepts a box the ray's LINE crosses behind its origin (entry < exit < 0,
// objects.cpp:353-376 has no `exit >= 0`), so it walks about as many nodes behind a ray as in front of it; every triangle
// there is then rejected by `t <= bias` (objects.cpp:227). Not visiting such a box changes no hit, no distance and no order
// among the remaining visits, PROVIDED no triangle inside it could be accepted — which rounding can break in one way only:
// a triangle whose plane passes within rounding noise of the origin gets a t = -pz/dz of arbitrary sign. Hence the margin:
// a box is dropped only if, on some axis, all of it lies behind the origin by more than delta = 1e-4 * (|p| + box extent)
// (rounding noise of pz is ~1e-7 of the same scale). Then for a triangle in it, either that axis is the one its projection
// drops — its plane is then more than delta / sqrt(3) from the origin on the far side, t comes out negative — or the axis
// is kept, and every point of the forward ray is outside the triangle's bounds by more than delta, so the barycentric
// test rejects it (its own rounding is ~1e-6 relative). In the ray's parameter the condition is exit < -delta_a * |1/d_a|;
// behind_bound returns the most negative of the three, one number per (ray, mesh). Rays with a degenerate axis take the
// other slab path and are not culled this way.
__device__ __forceinline__ float behind_bound(const ObjRec &o, const Ray3 &r, const V3 &drcp)
{
  if (qabs(r.d.x) < 1e-7f || qabs(r.d.y) < 1e-7f || qabs(r.d.z) < 1e-7f) return -QB_BIGFLOAT;
  const float dx = (qabs(r.p.x) + qmax(qabs(o.lo.x), qabs(o.hi.x))) * qabs(drcp.x);
  const float dy = (qabs(r.p.y) + qmax(qabs(o.lo.y), qabs(o.hi.y))) * qabs(drcp.y);
  const float dz = (qabs(r.p.z) + qmax(qabs(o.lo.z), qabs(o.hi.z))) * qabs(drcp.z);
  return -1e-4f * qmax(dx, qmax(dy, dz));
}

// IO concept:
//   uint32_t count() const                       number of rays
//   uint32_t *cursor()                           device work cursor
//   bool load(uint32_t idx, Ray3 &c, float &tmax)    world-space ray; false = padding entry (skipped, store not called)
//   void store(uint32_t idx, const HitRec &)     closest
//   void store_any(uint32_t idx, bool occluded)  any-hit
// QB_ANC_CACHE: objects are visited in scene-tree order, so the children of one group node follow each other (tower: three
// planes under one node, xmas: seven planes under one and seven meshes under another). The ray in the group's space is the
// same value for all of them — ToNodeCoords of the same ray through the same node — so it is kept and reused instead of
// recomputed per child: 9 -> 7 node transforms per ray on the tower, 29 -> 17 on xmas, bit for bit the same rays.
// The 7 words per lane (node id + ray) sit in QB_ANC_WORDS extra stack entries above the deepest entry a walk can reach
// (the host adds them to DScene::bvh_stack and sets DScene::anc_cache) — but only where they do not cost a resident block:
// on the tower the deeper stacks leave no room at 8 blocks / SM and the cache would lose more (75.3 ms per step at 7 blocks
// against 71.5) than it saves; on xmas it fits: 425.4 -> 415.2 ms.
#ifndef QB_ANC_CACHE
#define QB_ANC_CACHE 1
#endif
// Words of a lane's COLD state, kept in shared memory ([word][thread], thread-private: no synchronisation, no bank
// conflicts): what only the NEXT phase (a few times per ray) or a rare branch touches. With these 17 values in registers
// the closest-hit kernels needed 80 registers (6 blocks / SM); without them they fit 64 and run 8 blocks / SM — a third
// more warps to hide the node and triangle loads behind (long_scoreboard was the top stall).
enum {
  CW_C0PX = 0, CW_C0PY, CW_C0PZ, CW_C0DX, CW_C0DY, CW_C0DZ,        // the ray in the root node's space
  CW_HK, CW_HSLOT, CW_HTEXK, CW_HTEXSLOT, CW_HTEXT, CW_HLASTTRI,  // HitRec without z
  CW_ZENTRY, CW_ZSANE, CW_MEXIT,                                  // degenerate-axis cull of the current mesh
  CW_UVSLOT, CW_UVT,                                              // last accepted triangle with uv inside the current object
  CW_COUNT
};
// shared memory of one block of a traversal kernel, in words: stack[entries][QB_BLOCK], cold[CW_COUNT][QB_BLOCK], leaf scratch
__host__ __device__ inline size_t lanepool_smem_bytes(int entries) { return 4u * ((size_t) (entries + CW_COUNT) * QB_BLOCK + (QB_BLOCK / 32) * QB_COOP_SLOTS); }

template <bool ANY, class IO>
__device__ __forceinline__ void traverse_pool(const DScene &S, IO &io, uint32_t *__restrict__ s_dyn, const SceneTab &tab, TraceCounters &cnt)
{
  TabRef T = tab_ref(tab);
  const int lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const uint32_t n = io.count();
  uint32_t *stk = s_dyn + threadIdx.x;  // entry e at stk[e * QB_BLOCK]
#if QB_STACK_ASM
  // The same entries through 32-bit shared-space addresses: nvcc re-derives the shared window base for every access made
  // through the generic pointer (S2UR SR_CgaCtaId + 3 uniform ops before each STS / LDS: ~11 issue slots per node visit for
  // its push and pop); with the lane's base kept in one register an access is one IMAD + the STS / LDS.
  const uint32_t stk_a = (uint32_t) __cvta_generic_to_shared(stk);
#define STK_PUT(e, v) asm volatile("st.shared.u32 [%0], %1;" ::"r"(stk_a + (uint32_t) (e) * (QB_BLOCK * 4u)), "r"(v))
#define STK_GET(e, v) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(stk_a + (uint32_t) (e) * (QB_BLOCK * 4u)))
#else
#define STK_PUT(e, v) (stk[(e) * QB_BLOCK] = (v))
#define STK_GET(e, v) ((v) = stk[(e) * QB_BLOCK])
#endif
  uint32_t *cold = s_dyn + S.bvh_stack * QB_BLOCK + threadIdx.x;
  uint32_t *s_coop = s_dyn + (S.bvh_stack + CW_COUNT) * QB_BLOCK;
#if QB_STACK_ASM
  const uint32_t cold_a = smem_addr(cold);
#define CWGET(f) lds_u32(cold_a + (uint32_t) (f) * (QB_BLOCK * 4u))
#define CWSET(f, v) sts_u32(cold_a + (uint32_t) (f) * (QB_BLOCK * 4u), (v))
#else
#define CWGET(f) (cold[(f) * QB_BLOCK])
#define CWSET(f, v) (cold[(f) * QB_BLOCK] = (v))
#endif
#define CWF(f) __uint_as_float(CWGET(f))

  bool active = false, exhausted = false;
  uint32_t idx = 0;
  Ray3 r;
  V3 drcp = v3(0.f);
  bool gx = false, gy = false, gz = false;
  int cull_state = 0;     // 0 nothing dropped by the cull yet, 1 something dropped, 2 cull off (exact repeat of the mesh)
  int k = 0, sp = 0;
  uint32_t cur = QB_REF_EMPTY;
  float hz = 0.f;     // HitRec::z, the running hit distance (the rest of the record is cold)
  float bcull = -QB_BIGFLOAT;  // boxes the ray leaves before this (negative) parameter lie behind its origin: see behind_bound
  int obj_slot = -1;  // last accepted triangle inside the current object
  // cold: CW_ZENTRY = hz when the current mesh was entered, CW_ZSANE = hits of the current mesh beyond this distance can only
  // be rounding artefacts, CW_MEXIT = where the ray leaves the current mesh's box (bounds the degenerate-axis cull below)

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

    // weighted plurality (weights in quarters): the cheap phases may be given more than their head count
    const int w_inner = 4 * n_inner, w_leaf = (ANY ? QB_W_LEAF_ANY : QB_W_LEAF) * n_leaf, w_next = (ANY ? QB_W_NEXT_ANY : QB_W_NEXT) * n_next,
              w_fetch = (ANY ? QB_W_FETCH_ANY : QB_W_FETCH) * n_fetch;
    if (w_inner >= w_leaf && w_inner >= w_next && w_inner >= w_fetch) {
      // ---- INNER: node visits (both children's boxes); the near child becomes current, the far child is deferred.
      // Up to QB_INNER_REPS visits per vote while at least half of the lanes that started are still on inner nodes.
#pragma unroll 1
      for (int rep = 0; rep < QB_INNER_REPS; ++rep) {
        const bool go = active && cur != QB_REF_EMPTY && !(cur & QB_LEAF_BIT);
        if (rep > 0 && 2 * __popc(__ballot_sync(0xFFFFFFFFu, go)) < n_inner) break;
        if (go) {
          const DBvhNode *np = &S.bvh[cur];
          float4 na, nb, nc, nd4;
          ld_rec64(np, na, nb, nc, nd4);
          const uint2 ndr = make_uint2(__float_as_uint(nd4.x), __float_as_uint(nd4.y));
          QB_CNT(nodes);
          const float p00x = -(r.p.x - na.x) * drcp.x, p00y = -(r.p.y - na.y) * drcp.y, p00z = -(r.p.z - na.z) * drcp.z;
          const float p01x = -(r.p.x - na.w) * drcp.x, p01y = -(r.p.y - nb.x) * drcp.y, p01z = -(r.p.z - nb.y) * drcp.z;
          const float p10x = -(r.p.x - nb.z) * drcp.x, p10y = -(r.p.y - nb.w) * drcp.y, p10z = -(r.p.z - nc.x) * drcp.z;
          const float p11x = -(r.p.x - nc.y) * drcp.x, p11y = -(r.p.y - nc.z) * drcp.y, p11z = -(r.p.z - nc.w) * drcp.z;
          const float t_max = hz;
          float entry0, entry1, exit0, exit1;
          bool h0, h1;
          if (!(gx || gy || gz)) {
            // no degenerate axis: every p is finite or +-inf, never NaN, so the reference's MIN/MAX macros
            // (x < y ? x : y) and fminf/fmaxf pick the same value up to the sign of a zero, which no comparison sees
            entry0 = fmaxf(fminf(p00x, p01x), fmaxf(fminf(p00y, p01y), fminf(p00z, p01z)));
            entry1 = fmaxf(fminf(p10x, p11x), fmaxf(fminf(p10y, p11y), fminf(p10z, p11z)));
            exit0 = fminf(fmaxf(p00x, p01x), fminf(fmaxf(p00y, p01y), fmaxf(p00z, p01z)));
            exit1 = fminf(fmaxf(p10x, p11x), fminf(fmaxf(p10y, p11y), fmaxf(p10z, p11z)));
            // ... and boxes that lie behind the ray's origin by more than the margin of behind_bound are not visited
            h0 = (entry0 < t_max && entry0 < exit0) && !(exit0 < bcull);
            h1 = (entry1 < t_max && entry1 < exit1) && !(exit1 < bcull);
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
            const float t_far = qmin(t_max, CWF(CW_MEXIT));
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
            if (both) STK_PUT(sp, far_ref);
            sp += both ? 1 : 0;
            const bool pop = none && sp > 0;
            sp -= pop ? 1 : 0;
            uint32_t popped = QB_REF_EMPTY;
            if (pop) STK_GET(sp, popped);
            cur = none ? popped : near_ref;
          }
        }
      }
    } else if (w_leaf >= w_next && w_leaf >= w_fetch) {
      // ---- LEAF: the leaves' triangles, pooled over the warp -------------------------------------------------------------
      // A lane looping over its own leaf's 1..8 triangles kept the other lanes of the phase waiting for the longest leaf
      // (9 of 32 threads per instruction in the round-1 capture). The (ray, triangle) pairs of all leaf lanes are laid out
      // in one list instead and every lane of the warp tests one pair per pass — everything of IntersectTriangle that does
      // not depend on the running hit distance — with the owner's ray fetched by shuffle; the owner then applies
      // `hInfo.z > t` to its results in elements[] order, so the accepted triangle and t are what the serial loop gives.
#if !QB_COOP_LEAF
      if (s_leaf) {  // A/B: the round-1 per-lane loop
        const uint32_t base = cur & 0x0FFFFFFFu;
        const uint32_t count = ((cur >> 28) & 7u) + 1u;
        cur = sp > 0 ? stk[(--sp) * QB_BLOCK] : QB_REF_EMPTY;
        for (uint32_t i = 0; i < count; ++i) {
          const DTri *tp = &S.tris[base + i];
          float4 ta, tb, tc, td;
          ld_rec64(tp, ta, tb, tc, td);
          QB_CNT(tris);
          float t;
          if (tri_accept(r, hz, xyz(ta), xyz(tb), xyz(tc), v3(ta.w, tb.w, tc.w), td.x, __float_as_int(td.y), t)) {
            if (ANY) { io.store_any(idx, true); active = false; break; }
            hz = t;
            obj_slot = (int) (base + i);
            if (__float_as_uint(td.z) & 0x80000000u) { CWSET(CW_UVSLOT, (uint32_t) obj_slot); CWSET(CW_UVT, __float_as_uint(t)); }
          }
        }
      }
#else
      {
        const uint32_t lbase = cur & 0x0FFFFFFFu;
        uint32_t count = s_leaf ? ((cur >> 28) & 7u) + 1u : 0u;
        uint32_t incl = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
          if (lane >= o) incl += v;
        }
        if (incl > QB_COOP_SLOTS) count = 0u;  // does not fit this round: the lane keeps its leaf for the next LEAF phase
        const uint32_t excl = incl - count;
        const uint32_t total = min(__reduce_max_sync(0xFFFFFFFFu, count ? incl : 0u), (uint32_t) QB_COOP_SLOTS);
        uint32_t *scr = s_coop + (threadIdx.x >> 5) * QB_COOP_SLOTS;
#if QB_STACK_ASM
        const uint32_t scr_a = smem_addr(scr);
#define SCR_PUT(i, v) sts_u32(scr_a + 4u * (uint32_t) (i), (v))
#define SCR_GET(i) lds_u32(scr_a + 4u * (uint32_t) (i))
#else
#define SCR_PUT(i, v) (scr[(i)] = (v))
#define SCR_GET(i) (scr[(i)])
#endif
        for (uint32_t i = 0; i < count; ++i) SCR_PUT(excl + i, ((uint32_t) lane << 3) | i);
        __syncwarp();
        for (uint32_t q0 = 0; q0 < total; q0 += 32u) {
          const uint32_t q = q0 + (uint32_t) lane;
          const bool have = q < total;
          uint32_t code = 0u;
          if (have) code = SCR_GET(q);
          const int o = (int) (code >> 3);
          const uint32_t tb = __shfl_sync(0xFFFFFFFFu, lbase, o);
          Ray3 rr;
          rr.p.x = __shfl_sync(0xFFFFFFFFu, r.p.x, o); rr.p.y = __shfl_sync(0xFFFFFFFFu, r.p.y, o); rr.p.z = __shfl_sync(0xFFFFFFFFu, r.p.z, o);
          rr.d.x = __shfl_sync(0xFFFFFFFFu, r.d.x, o); rr.d.y = __shfl_sync(0xFFFFFFFFu, r.d.y, o); rr.d.z = __shfl_sync(0xFFFFFFFFu, r.d.z, o);
          if (have) {
            const DTri *tp = &S.tris[tb + (code & 7u)];
            float4 ta, tb4, tc, td;
            ld_rec64(tp, ta, tb4, tc, td);
            QB_CNT(tris);
            float t;
            const bool ok = tri_candidate(rr, xyz(ta), xyz(tb4), xyz(tc), v3(ta.w, tb4.w, tc.w), td.x, __float_as_int(td.y), t);
            // accepted candidates have t > bias > 0; bit 0 of a rejected entry's word is never looked at
            SCR_PUT(q, ok ? __float_as_uint(t) : 0x80000000u);
          }
        }
        __syncwarp();
        if (count) {
          cur = QB_REF_EMPTY;
          if (sp > 0) { --sp; STK_GET(sp, cur); }
          for (uint32_t i = 0; i < count; ++i) {
            const uint32_t w = SCR_GET(excl + i);
            const float t = __uint_as_float(w);
            if (!(w & 0x80000000u) && hz > t) {
              if (ANY) { io.store_any(idx, true); active = false; break; }
              hz = t;
              obj_slot = (int) (lbase + i);
              if (__ldg(&S.tri_face[lbase + i]) & 0x80000000u) { CWSET(CW_UVSLOT, (uint32_t) obj_slot); CWSET(CW_UVT, __float_as_uint(t)); }
            }
          }
        }
        __syncwarp();  // the scratch is rewritten by the next LEAF phase
      }
#endif
    } else if (w_next >= w_fetch) {
      // ---- NEXT: the object just traversed reports its result, then the next object of the list is set up ----------
      // Up to QB_NEXT_REPS objects per vote while at least half of the lanes that started are still between objects (a
      // plane / sphere test or a missed mesh box leaves the lane in this state): a ray crosses the whole object list in a
      // few NEXT phases instead of one per object.
#pragma unroll 1
      for (int rep = 0; rep < QB_NEXT_REPS; ++rep) {
      const bool go_next = active && cur == QB_REF_EMPTY;
      if (rep > 0 && 2 * __popc(__ballot_sync(0xFFFFFFFFu, go_next)) < n_next) break;
      bool redo = false, handed_over = false;
      if (go_next && k > 0 && cull_state == 1) {
        // the cull dropped boxes and the mesh produced no hit inside its own box: repeat it exactly as the reference walks it
        cull_state = 0;
        // closest-hit rays of the wavefront hand the check to k_redo_detect / k_redo_exact (trace_kernels.cu) and go on: a
        // single lane repeating a whole mesh costs up to ~30 ms, and some sample ranges hold a few dozen such rays
        const float z_entry = CWF(CW_ZENTRY);
        if (!(hz <= CWF(CW_ZSANE))) {
          // any-hit rays (a directional light's shadow ray has no far end: every unoccluded one lands here when its direction
          // has a degenerate axis in the mesh's frame) hand the WHOLE ray to k_any_exact, which walks it as the reference
          // does with a warp per ray; the lane drops it. A full queue, and the closest-hit rays of the ray-buffer / redo
          // passes whose queue is full, repeat the mesh here.
          if (ANY && io.defer(idx, k - 1, z_entry)) {
            active = false;
            handed_over = true;
          } else if (ANY || !io.defer(idx, k - 1, z_entry)) {
            redo = true;
            cull_state = 2;
            hz = z_entry;
            obj_slot = -1; CWSET(CW_UVSLOT, 0xFFFFFFFFu); sp = 0;
            cur = __float_as_uint(obj_rec(S, T, k - 1).lo.w);
          }
        }
      }
      if (go_next && !redo && !handed_over) {
        if (!ANY && k > 0 && obj_slot >= 0) {
          CWSET(CW_HK, (uint32_t) (k - 1)); CWSET(CW_HSLOT, (uint32_t) obj_slot); CWSET(CW_HLASTTRI, (uint32_t) obj_slot);
          const uint32_t uv_slot = CWGET(CW_UVSLOT);
          if ((int) uv_slot >= 0) { CWSET(CW_HTEXK, (uint32_t) (k - 1)); CWSET(CW_HTEXSLOT, uv_slot); CWSET(CW_HTEXT, CWGET(CW_UVT)); }
        }
        obj_slot = -1;
        if (k == S.num_obj) {
          if (ANY) {
            io.store_any(idx, false);
          } else {
            HitRec h;
            h.z = hz; h.k = (int) CWGET(CW_HK); h.slot = (int) CWGET(CW_HSLOT); h.tex_k = (int) CWGET(CW_HTEXK); h.tex_slot = (int) CWGET(CW_HTEXSLOT);
            h.tex_t = CWF(CW_HTEXT); h.last_tri = (int) CWGET(CW_HLASTTRI);
            io.store(idx, h);
          }
          active = false;
        } else {
          const ObjRec o = obj_rec(S, T, k);
          Ray3 c0;
          c0.p = v3(CWF(CW_C0PX), CWF(CW_C0PY), CWF(CW_C0PZ));
          c0.d = v3(CWF(CW_C0DX), CWF(CW_C0DY), CWF(CW_C0DZ));
#if QB_ANC_CACHE
          if (S.anc_cache && k < QB_TAB_OBJS && tab_chain_ok(T) && o.id.w - o.id.z >= 2) {
            const int anc = tab_chain(T, o.id.z);
            uint32_t *aw = stk + (S.bvh_stack - QB_ANC_WORDS) * QB_BLOCK;  // word w of this lane at aw[w * QB_BLOCK]
            if ((int) aw[0] == anc) {
              r.p = v3(__uint_as_float(aw[1 * QB_BLOCK]), __uint_as_float(aw[2 * QB_BLOCK]), __uint_as_float(aw[3 * QB_BLOCK]));
              r.d = v3(__uint_as_float(aw[4 * QB_BLOCK]), __uint_as_float(aw[5 * QB_BLOCK]), __uint_as_float(aw[6 * QB_BLOCK]));
            } else {
              r = to_node_tab(S, T, anc, c0);
              aw[0] = (uint32_t) anc;
              aw[1 * QB_BLOCK] = __float_as_uint(r.p.x); aw[2 * QB_BLOCK] = __float_as_uint(r.p.y); aw[3 * QB_BLOCK] = __float_as_uint(r.p.z);
              aw[4 * QB_BLOCK] = __float_as_uint(r.d.x); aw[5 * QB_BLOCK] = __float_as_uint(r.d.y); aw[6 * QB_BLOCK] = __float_as_uint(r.d.z);
            }
            for (int c = o.id.z + 1; c < o.id.w; ++c) r = to_node_tab(S, T, tab_chain(T, c), r);
          } else
#endif
          r = chain_ray_tab(S, T, o, k, c0);
          const int type = o.id.x;
          QB_CNT(objects);
          ++k;
          CWSET(CW_UVSLOT, 0xFFFFFFFFu);
          if (type == QG_OBJ_MESH) {
            const uint32_t root_ref = __float_as_uint(o.lo.w);
            float mesh_exit;
            if (root_ref != QB_NO_FACES && mesh_box_hit_q(o, r, hz, mesh_exit, drcp) && !(mesh_exit < (bcull = behind_bound(o, r, drcp)))) {
              cur = root_ref;
              cull_state = io.exact(idx, k - 1) ? 2 : 0;  // k was advanced above: k - 1 is this object
              CWSET(CW_ZENTRY, __float_as_uint(hz));
              CWSET(CW_ZSANE, __float_as_uint(qmax(mesh_exit * 1.001f, o.hi.w)));
              CWSET(CW_MEXIT, __float_as_uint(mesh_exit));
              gx = qabs(r.d.x) < 1e-7f; gy = qabs(r.d.y) < 1e-7f; gz = qabs(r.d.z) < 1e-7f;
            }
          } else {
            const float t = type == QG_OBJ_SPHERE ? sphere_t(r, hz) : plane_t(r, hz);
            if (t > 0.f) {
              if (ANY) {
                io.store_any(idx, true);
                active = false;
              } else {
                hz = t;
                CWSET(CW_HK, (uint32_t) (k - 1)); CWSET(CW_HSLOT, 0xFFFFFFFFu);
                CWSET(CW_HTEXK, (uint32_t) (k - 1)); CWSET(CW_HTEXSLOT, 0xFFFFFFFFu); CWSET(CW_HTEXT, __float_as_uint(t));  // sphere / plane always write uvw
              }
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
          const Ray3 c0 = to_node_tab(S, T, 0, c);
          CWSET(CW_C0PX, __float_as_uint(c0.p.x)); CWSET(CW_C0PY, __float_as_uint(c0.p.y)); CWSET(CW_C0PZ, __float_as_uint(c0.p.z));
          CWSET(CW_C0DX, __float_as_uint(c0.d.x)); CWSET(CW_C0DY, __float_as_uint(c0.d.y)); CWSET(CW_C0DZ, __float_as_uint(c0.d.z));
          hz = tmax;
          if (!ANY) {
            CWSET(CW_HK, 0xFFFFFFFFu); CWSET(CW_HSLOT, 0xFFFFFFFFu); CWSET(CW_HTEXK, 0xFFFFFFFFu); CWSET(CW_HTEXSLOT, 0xFFFFFFFFu);
            CWSET(CW_HTEXT, 0u); CWSET(CW_HLASTTRI, 0xFFFFFFFFu);
          }
          CWSET(CW_UVSLOT, 0xFFFFFFFFu);
#if QB_ANC_CACHE
          if (S.anc_cache) stk[(S.bvh_stack - QB_ANC_WORDS) * QB_BLOCK] = 0xFFFFFFFFu;
#endif
          k = 0; sp = 0; cur = QB_REF_EMPTY; obj_slot = -1;
          active = true;
        }
      }
    }
  }
#undef CWGET
#undef CWSET
#undef CWF
}

// Attributes of a finished closest-hit ray (what Scene::TraceNodeNormal leaves in DiffHitInfo::c): the tail of
// traversal, run by full warps afterwards. c/x/y: world-space rays; with_diff = false treats x and y as absent (duvw = 0),
// the form used for secondary rays. stale_tex: reproduce the reference's carry-over of uvw/duvw/hasTexture from an earlier,
// farther hit when the closest hit is a triangle without texture coordinates (IntersectTriangle leaves those fields untouched).
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
