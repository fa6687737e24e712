// trace_kernels.cu — ray generation and traversal kernels (sm_100a). Compiled with --fmad=false (see trace.cuh).
//
//   k_wf_trace     closest-hit over the wave's path queue; at depth 0 it builds the primary DiffRay of
//                  Renderer::PixelRender (src/renderers/renderer.cpp:312-326) in registers instead of reading a ray buffer
//   k_wf_shadow    any-hit over the shadow queue (GenLight::Shadow, src/lights/lights.cpp:39-48); unoccluded rays
//                  add their pre-multiplied contribution to the fp32 accumulation buffer
//   k_trace_*_rays ray-buffer entry points of the C ABI (qg_trace_closest / qg_trace_any / qg_primary_rays)
//
// Persistent kernels: grid = SMs x resident blocks; each warp pulls 32 consecutive rays from a device cursor, so
// long rays do not hold back a statically assigned tail and no host sync is needed between bounces.
#include "traverse.cuh"
#include "blockpool.cuh"
#include "anyexact.cuh"
#include "wavefront.h"
#include "wave_rays.cuh"
#include <algorithm>
#include <map>
#include <mutex>
#include <set>
#include <tuple>
#include <utility>

#ifndef QB_BLOCKPOOL
#define QB_BLOCKPOOL 0  // 0: per-warp lane pool (traverse.cuh); 1: block-level ray pool (blockpool.cuh), measured 20 % slower, kept for A/B
#endif

namespace qb {

#if QB_BLOCKPOOL
#define QB_TRACE_THREADS QB_BP_THREADS
#define QB_TRACE_BOUNDS __launch_bounds__(QB_BP_THREADS, QB_BP_BLOCKS)
#define QB_TRACE_BOUNDS_ANY __launch_bounds__(QB_BP_THREADS, QB_BP_BLOCKS)
// the traversal of one kernel: scene table, then the block pool over dynamic shared memory
#define QB_TRAVERSE(ANYHIT, io, cnt)                          \
  extern __shared__ __align__(16) uint32_t s_pool[];          \
  __shared__ SceneTab s_tab;                                  \
  load_scene_tab(S, s_tab);                                   \
  traverse_block<ANYHIT>(S, io, s_pool, S.bvh_stack, s_tab, cnt)
#else
#define QB_TRACE_THREADS QB_BLOCK
#define QB_TRACE_BOUNDS __launch_bounds__(QB_BLOCK, QB_MIN_BLOCKS)
#define QB_TRACE_BOUNDS_ANY __launch_bounds__(QB_BLOCK, QB_MIN_BLOCKS_ANY)
// dynamic shared memory: the lanes' stacks (as deep as the scene's deepest BVH), their cold state, the leaf scratch
#define QB_TRAVERSE(ANYHIT, io, cnt)                          \
  extern __shared__ __align__(16) uint32_t s_dyn[];           \
  __shared__ SceneTab s_tab;                                  \
  load_scene_tab(S, s_tab);                                   \
  traverse_pool<ANYHIT>(S, io, s_dyn, s_tab, cnt)
#endif

// counting build only: stats[2..4] = BVH nodes / triangles / objects visited by closest-hit rays, [5..7] by any-hit rays
__device__ __forceinline__ void flush_counters(unsigned long long *stats, const TraceCounters &cnt, int base = 2)
{
#ifdef QB_COUNT
  atomicAdd(&stats[base + 0], cnt.nodes);
  atomicAdd(&stats[base + 1], cnt.tris);
  atomicAdd(&stats[base + 2], cnt.objects);
#endif
}

// ---- IO policies of traverse_pool ------------------------------------------------------------------------------------
// QB_STREAM_CS: the wave state is read and written once per launch — mark it evict-first (ld/st.global.cs) so that it does
// not push the scene's records out of L2
#ifndef QB_STREAM_CS
#define QB_STREAM_CS 0
#endif
__device__ __forceinline__ float4 ld_stream(const float4 *p)
{
#if QB_STREAM_CS
  return __ldcs(p);
#else
  return *p;
#endif
}
__device__ __forceinline__ void st_stream(float4 *p, const float4 v)
{
#if QB_STREAM_CS
  __stcs(p, v);
#else
  *p = v;
#endif
}
// compact hit record of the wavefront (wave_rays.cuh): r0 = z, k, slot, tex_slot ; r1 = tex_t, tex_k, last_tri, -
__device__ __forceinline__ void put_rec(float4 *r0, float4 *r1, uint32_t idx, const HitRec &h)
{
  st_stream(&r0[idx], f4(h.z, __int_as_float(h.k), __int_as_float(h.slot), __int_as_float(h.tex_slot)));
  st_stream(&r1[idx], f4(h.tex_t, __int_as_float(h.tex_k), __int_as_float(h.last_tri), 0.f));
}
// the same without touching r1.w: the cull-check pass keeps its flag bits there while a ray may be walked more than once
__device__ __forceinline__ void put_rec_keep_w(float4 *r0, float4 *r1, uint32_t idx, const HitRec &h)
{
  r0[idx] = f4(h.z, __int_as_float(h.k), __int_as_float(h.slot), __int_as_float(h.tex_slot));
  float *p = (float *) &r1[idx];
  p[0] = h.tex_t; p[1] = __int_as_float(h.tex_k); p[2] = __int_as_float(h.last_tri);
}
struct RedoQueue {
  uint4 *q;          // {ray index, object, z at mesh entry (bits), flag}
  uint32_t cap;
  uint32_t *count;   // entries pushed (may run past cap: clamp)
  uint32_t *fetch;   // work cursor of the exact pass
};
__device__ __forceinline__ bool redo_push(const RedoQueue &Q, uint32_t idx, int k, float z_entry)
{
  if (Q.cap == 0u) return false;
  const uint32_t q = atomicAdd(Q.count, 1u);
  if (q >= Q.cap) return false;
  Q.q[q] = make_uint4(idx, (uint32_t) k, __float_as_uint(z_entry), 0u);
  return true;
}
template <bool PRIMARY>
struct WfClosestIO {
  const WfBuffers &B;
  const WfParams &P;
  uint32_t n;
  __device__ uint32_t count() const { return n; }
  __device__ uint32_t *cursor() { return &B.counters[QB_CNT_FETCH_TRACE + P.depth]; }
  __device__ uint32_t map(uint32_t q) const { return q; }
  __device__ bool exact(uint32_t, int) const { return false; }  // true: walk this mesh with the degenerate-axis cull off
  // hand a (ray, mesh) whose culled walk dropped boxes and found nothing to the redo kernels; false when the queue is full
  __device__ bool defer(uint32_t idx, int k, float z_entry) const
  {
    return redo_push(RedoQueue{B.redo, B.redo_capacity, &B.counters[QB_CNT_REDO + P.depth], nullptr}, idx, k, z_entry);
  }
  __device__ bool load(uint32_t idx, Ray3 &c, float &tmax)
  {
    tmax = QB_BIGFLOAT;
    if (PRIMARY) {
      int i, j, s, slot;
      if (!wf_decode(P, idx, i, j, s, slot)) {
        B.h0[idx] = f4(QB_BIGFLOAT, __int_as_float(-2), 0.f, 0.f);  // padding lane of an edge tile
        return false;
      }
      Ray3 x, y;
      primary_ray(P, i, j, s, c, x, y);
    } else {
      const float4 a = ld_stream(&B.ps0[P.depth & 1][idx]), b = ld_stream(&B.ps1[P.depth & 1][idx]);
      if (__float_as_int(b.z) < 0) {  // dead entry (shade_kernels.cu, OutRes): report a miss, trace nothing
        HitRec m;
        m.z = QB_BIGFLOAT; m.k = -1; m.slot = -1; m.tex_k = -1; m.tex_slot = -1; m.tex_t = 0.f; m.last_tri = -1;
        put_rec(B.h1, B.h2, idx, m);
        return false;
      }
      c.p = v3(a.x, a.y, a.z);
      c.d = v3(a.w, b.x, b.y);
    }
    return true;
  }
  __device__ void store(uint32_t idx, const HitRec &h) { put_rec(B.h1, B.h2, idx, h); }  // h1/h2 double as the record until finish
  __device__ void store_any(uint32_t, bool) {}
  // cull check: bit k of the record's spare word = "object k of this ray must be walked with the cull off" (k_redo_detect)
  __device__ void store_redo(uint32_t idx, const HitRec &h) { put_rec_keep_w(B.h1, B.h2, idx, h); }
  __device__ void flag(uint32_t idx, int k) const { if (k < 32) atomicOr((uint32_t *) &B.h2[idx].w, 1u << k); }
  __device__ uint32_t flags(uint32_t idx) const { return __float_as_uint(B.h2[idx].w); }
};

struct WfShadowIO {
  const WfBuffers &B;
  const WfParams &P;
  uint32_t n;
  __device__ uint32_t count() const { return n; }
  __device__ uint32_t *cursor() { return &B.counters[QB_CNT_FETCH_SHADOW + P.depth]; }
  __device__ uint32_t map(uint32_t q) const { return q; }                 // queue position -> ray index
  __device__ bool exact(uint32_t, int) const { return false; }  // true: walk this mesh with the degenerate-axis cull off
  __device__ AnyQueue any_queue() const
  {
    return AnyQueue{B.anyq, B.anyq_capacity, &B.counters[QB_CNT_ANYQ + P.depth], &B.counters[QB_CNT_FETCH_ANYQ + P.depth]};
  }
  // true: the whole ray goes to k_any_exact (a warp walks it as the reference does); false: the lane repeats the mesh itself
  __device__ bool defer(uint32_t idx, int, float) const { return any_push(any_queue(), idx); }
  __device__ bool load(uint32_t idx, Ray3 &c, float &tmax)
  {
    const float4 a = ld_stream(&B.s0[idx]), b = ld_stream(&B.s1[idx]);
    c.p = v3(a.x, a.y, a.z);
    c.d = v3(a.w, b.x, b.y);
    tmax = b.z;
    return tmax >= 0.f;  // tmax -1: a dead entry (shade_kernels.cu, OutRes)
  }
  __device__ void store(uint32_t, const HitRec &) {}
  __device__ void store_any(uint32_t idx, bool occluded)
  {
    if (occluded) return;
    const float4 c = B.s2[idx];
    float *acc = P.accum + 4ll * (long long) __float_as_int(B.s1[idx].w);
    atomicAdd(acc + 0, c.x);
    atomicAdd(acc + 1, c.y);
    atomicAdd(acc + 2, c.z);
  }
};

// PRIMARY = depth 0 (rays generated in registers) / depth >= 1 (rays read from the path state): two instantiations.
template <bool PRIMARY>
__global__ void QB_TRACE_BOUNDS k_wf_trace(const DScene S, const WfBuffers B, const WfParams P)
{
  TraceCounters cnt = {0, 0, 0};
  WfClosestIO<PRIMARY> io{B, P, PRIMARY ? P.n_paths : min(B.counters[QB_CNT_PATHS + P.depth], B.capacity)};
  QB_TRAVERSE(false, io, cnt);
  flush_counters(B.stats, cnt);
}

// ---- deferred check of the degenerate-axis cull ----------------------------------------------------------------------------
// trace(d) queued the (ray, mesh) pairs whose culled walk dropped boxes and found nothing (traverse.cuh). The reference walks
// every box such a ray's two regular axes cross and can accept a rounding-noise hit there (t ~ 1e6 on a triangle the ray
// never touches). k_redo_detect decides with the whole GPU whether ANY triangle of the mesh accepts the ray — a superset of
// what the reference tests, so "none" proves the stored record exact (the usual outcome). A pair with an accepting triangle
// is re-traced by k_redo_exact with the cull off for the flagged meshes, i.e. as the reference walks them, and its record
// is overwritten.
template <class IO> __device__ __forceinline__ void redo_detect(const DScene &S, IO &io, const RedoQueue &Q)
{
  const uint32_t n = min(*Q.count, Q.cap);
  const DNode *root = &S.nodes[0];
  for (uint32_t e = 0; e < n; ++e) {
    const uint4 ent = Q.q[e];
    Ray3 c;
    float tmax;
    if (!io.load(ent.x, c, tmax)) continue;
    const int k = (int) ent.y;
    const Ray3 r = chain_ray(S, k, to_node(root, c));
    const DMesh *m = &S.meshes[__ldg(&S.nodes[__ldg(&S.obj_node[k])].mesh)];
    const uint32_t first = __ldg(&m->face_base), count = __ldg(&m->num_faces);  // triangle slots of a mesh = its face range
    const float z_entry = __uint_as_float(ent.z);
    // Only a hit beyond the mesh box matters: a hit inside it lies in boxes the ray reaches, which the cull never drops, so
    // the culled walk has tested it already (or the reference's own strict slab test skips its box, e.g. a flat one)
    float mesh_exit = 0.f;
    V3 drcp;
    if (!mesh_box_hit(m, r, z_entry, mesh_exit, drcp)) continue;
    const float t_box = mesh_exit * 1.001f;
    bool any = false;
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < count; t += gridDim.x * blockDim.x) {
      const DTri *tp = &S.tris[first + t];
      const float4 ta = __ldg(&tp->a), tb = __ldg(&tp->b), tc = __ldg(&tp->c), td = __ldg(&tp->d);
      float tt;
      if (tri_accept(r, z_entry, xyz(ta), xyz(tb), xyz(tc), v3(ta.w, tb.w, tc.w), td.x, __float_as_int(td.y), tt) && tt > t_box) any = true;
    }
    if (any) {
      Q.q[e].w = 1u;
      io.flag(ent.x, k);
    }
  }
}
// the flagged pairs' rays through the same traversal core with the cull off; records overwritten through the base IO
template <class IO>
struct RedoIO {
  IO &base;
  RedoQueue Q;
  uint32_t n;
  __device__ uint32_t count() const { return n; }
  __device__ uint32_t *cursor() { return Q.fetch; }
  __device__ uint32_t map(uint32_t q) const { return Q.q[q].w ? Q.q[q].x : 0xFFFFFFFFu; }
  // The cull is off only for the (ray, mesh) pairs k_redo_detect flagged: bit k of the ray's record word, the queue itself
  // for objects beyond 32. Every other pair that drops boxes and finds nothing in this pass did so in the first pass too
  // (h.z can only be smaller now) and was cleared there against a larger z, so it needs no second check: defer() swallows
  // it — unless the first pass overflowed the queue: pairs that did not fit were repeated exactly inside their lane then,
  // and are again now (defer() false), so an overflow costs time, never exactness.
  __device__ bool exact(uint32_t idx, int k) const
  {
    if (k < 32) return (base.flags(idx) >> k) & 1u;
    for (uint32_t e = 0; e < n; ++e) {
      const uint4 ent = Q.q[e];
      if (ent.x == idx && (int) ent.y == k && ent.w) return true;
    }
    return false;
  }
  __device__ bool defer(uint32_t, int, float) const { return *Q.count <= Q.cap; }
  __device__ bool load(uint32_t idx, Ray3 &c, float &tmax) { return idx != 0xFFFFFFFFu && base.load(idx, c, tmax); }
  __device__ void store(uint32_t idx, const HitRec &h) { base.store_redo(idx, h); }
  __device__ void store_any(uint32_t, bool) {}
};
__device__ __forceinline__ RedoQueue wf_redo_queue(const WfBuffers &B, const WfParams &P)
{
  return RedoQueue{B.redo, B.redo_capacity, &B.counters[QB_CNT_REDO + P.depth], &B.counters[QB_CNT_FETCH_REDO + P.depth]};
}
template <bool PRIMARY>
__global__ void __launch_bounds__(QB_BLOCK) k_redo_detect(const DScene S, const WfBuffers B, const WfParams P)
{
  if (B.counters[QB_CNT_REDO + P.depth] == 0u) return;
  WfClosestIO<PRIMARY> io{B, P, 0u};
  redo_detect(S, io, wf_redo_queue(B, P));
}
template <bool PRIMARY>
__global__ void QB_TRACE_BOUNDS k_redo_exact(const DScene S, const WfBuffers B, const WfParams P)
{
  const uint32_t n = min(B.counters[QB_CNT_REDO + P.depth], B.redo_capacity);
  if (n == 0u) return;
  TraceCounters cnt = {0, 0, 0};
  WfClosestIO<PRIMARY> base{B, P, 0u};
  RedoIO<WfClosestIO<PRIMARY>> io{base, wf_redo_queue(B, P), n};
  QB_TRAVERSE(false, io, cnt);
  flush_counters(B.stats, cnt);
}

// hit attributes of the wave's rays from the compact records, one thread per ray (full warps): h0..h3 for shade
#ifndef QB_FINISH_BLOCKS
#define QB_FINISH_BLOCKS 8
#endif
template <bool PRIMARY>
__global__ void __launch_bounds__(QB_BLOCK, QB_FINISH_BLOCKS) k_wf_finish(const DScene S, const WfBuffers B, const WfParams P)
{
  const uint32_t n = PRIMARY ? P.n_paths : min(B.counters[QB_CNT_PATHS + P.depth], B.capacity);
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
    Ray3 c, x, y;
    bool with_diff = PRIMARY;
    if (PRIMARY) {
      int i, j, s, slot;
      if (!wf_decode(P, idx, i, j, s, slot)) continue;  // h0 already marks the padding lane
      primary_ray(P, i, j, s, c, x, y);
    } else {
      const float4 a = B.ps0[P.depth & 1][idx], b = B.ps1[P.depth & 1][idx];
      c.p = v3(a.x, a.y, a.z);
      c.d = v3(a.w, b.x, b.y);
      x = c;
      y = c;
      // Whitted-style classes: the secondary ray is DiffRay(p, dir, px, dxdir, py, dydir) (MtlBlinn_Basic.cpp:131,150)
      if (P.carry_diff && (__float_as_uint(B.ps2[P.depth & 1][idx].w) & QB_FLAG_HAS_DIFF)) {
        const float4 e = B.ps4[P.depth & 1][idx], f = B.ps5[P.depth & 1][idx], g = B.ps6[P.depth & 1][idx];
        x.p = v3(e.x, e.y, e.z);
        x.d = v3(e.w, f.x, f.y);
        y.p = v3(f.z, f.w, g.x);
        y.d = v3(g.y, g.z, g.w);
        with_diff = true;
      }
    }
    const HitRec rec = get_rec(B.h1, B.h2, idx);
    if (rec.k < 0) {
      B.h0[idx] = f4(QB_BIGFLOAT, __int_as_float(-1), 0.f, 0.f);
      continue;
    }
    HitAttr h;
    float txy[2];
    finish_hit(S, rec, c, x, y, with_diff, P.stale_tex != 0, h, P.carry_diff ? txy : nullptr);
    if (P.carry_diff) B.h4[idx] = f4(txy[0], txy[1], 0.f, 0.f);
    const int flags = (h.front ? 1 : 0) | (h.has_texture ? 2 : 0);
    B.h0[idx] = f4(rec.z, __int_as_float(S.obj_node[rec.k]), __int_as_float(h.mtl_id), __int_as_float(flags));
    B.h1[idx] = f4(h.p.x, h.p.y, h.p.z, h.N.x);
    B.h2[idx] = f4(h.N.y, h.N.z, h.uvw.x, h.uvw.y);
    B.h3[idx] = f4(h.duvw0.x, h.duvw0.y, h.duvw1.x, h.duvw1.y);
  }
}

__global__ void QB_TRACE_BOUNDS_ANY k_wf_shadow(const DScene S, const WfBuffers B, const WfParams P)
{
  TraceCounters cnt = {0, 0, 0};
  WfShadowIO io{B, P, min(B.counters[QB_CNT_SHADOW + P.depth], B.shadow_capacity)};
  QB_TRAVERSE(true, io, cnt);
  flush_counters(B.stats, cnt, 5);
}

// the any-hit rays k_wf_shadow handed over (anyexact.cuh): a warp per ray, the reference's own walk
__global__ void __launch_bounds__(QB_ANYX_WARPS * 32) k_any_exact_shadow(const DScene S, const WfBuffers B, const WfParams P)
{
  __shared__ uint32_t s_stack[QB_ANYX_WARPS * QB_ANYX_STACK];
  if (B.counters[QB_CNT_ANYQ + P.depth] == 0u) return;
  TraceCounters cnt = {0, 0, 0};
  WfShadowIO io{B, P, 0u};
  any_exact_pass(S, io, io.any_queue(), s_stack, cnt);
  flush_counters(B.stats, cnt, 5);
}

// ---- area lights: PointLight / SpotLight with size > 0.01 (src/lights/lights.cpp:50-65,84-101) -----------------------------
// The reference estimates such a light's visibility with 16 shadow rays toward `position + UniformBall(size)`, and 64 when
// the running estimate ever lies strictly between 0 and 1. Sample s of a group is a pure function of (group key, s), so
// all 64 are traced here (any-hit, results as one bit each) and the resolve step replays the reference's loop over as many
// of them as it would have drawn — the unused ones change nothing.
struct WfAreaIO {
  const DScene &S;
  const WfBuffers &B;
  const WfParams &P;
  uint32_t n;  // QB_AREA_SPP_MAX * groups
  __device__ uint32_t count() const { return n; }
  __device__ uint32_t *cursor() { return &B.counters[QB_CNT_FETCH_AREA + P.depth]; }
  __device__ uint32_t map(uint32_t q) const { return q; }                 // queue position -> ray index
  __device__ bool exact(uint32_t, int) const { return false; }  // true: walk this mesh with the degenerate-axis cull off
  __device__ AnyQueue any_queue() const
  {
    return AnyQueue{B.anyq, B.anyq_capacity, &B.counters[QB_CNT_ANYQ_AREA + P.depth], &B.counters[QB_CNT_FETCH_ANYQ_AREA + P.depth]};
  }
  __device__ bool defer(uint32_t idx, int, float) const { return any_push(any_queue(), idx); }
  __device__ bool load(uint32_t idx, Ray3 &c, float &tmax)
  {
    const uint32_t g = idx / QB_AREA_SPP_MAX;
    const int s = (int) (idx % QB_AREA_SPP_MAX);
    const float4 a = B.a0[g];
    const V3 p = v3(a.x, a.y, a.z);
    const V3 dir = wf_area_dir(S.lights[__float_as_int(B.a1[g].x)], p, __float_as_uint(a.w), s);
    c.p = p;
    c.d = normalize(dir);  // Ray ray(p, dir); ray.Normalize(); Shadow(ray, length(dir))
    tmax = length(dir);
    return true;
  }
  __device__ void store(uint32_t, const HitRec &) {}
  __device__ void store_any(uint32_t idx, bool occluded)
  {
    if (occluded) return;
    const uint32_t g = idx / QB_AREA_SPP_MAX, s = idx % QB_AREA_SPP_MAX;
    atomicOr(&B.amask[2u * g + (s >> 5)], 1u << (s & 31u));
  }
};

__global__ void QB_TRACE_BOUNDS_ANY k_wf_area_shadow(const DScene S, const WfBuffers B, const WfParams P)
{
  TraceCounters cnt = {0, 0, 0};
  WfAreaIO io{S, B, P, min(B.counters[QB_CNT_AREA + P.depth], B.area_capacity) * (uint32_t) QB_AREA_SPP_MAX};
  QB_TRAVERSE(true, io, cnt);
  flush_counters(B.stats, cnt, 5);
}

__global__ void __launch_bounds__(QB_ANYX_WARPS * 32) k_any_exact_area(const DScene S, const WfBuffers B, const WfParams P)
{
  __shared__ uint32_t s_stack[QB_ANYX_WARPS * QB_ANYX_STACK];
  if (B.counters[QB_CNT_ANYQ_AREA + P.depth] == 0u) return;
  TraceCounters cnt = {0, 0, 0};
  WfAreaIO io{S, B, P, 0u};
  any_exact_pass(S, io, io.any_queue(), s_stack, cnt);
  flush_counters(B.stats, cnt, 5);
}

// PointLight::Illuminate's loop (lights.cpp:53-64) over the traced samples, one thread per group
__global__ void __launch_bounds__(QB_BLOCK) k_wf_area_resolve(const DScene S, const WfBuffers B, const WfParams P)
{
  const uint32_t n = min(B.counters[QB_CNT_AREA + P.depth], B.area_capacity);
  for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < n; g += gridDim.x * blockDim.x) {
    const float4 a = B.a0[g], b = B.a1[g], c = B.a2[g];
    const V3 p = v3(a.x, a.y, a.z);
    const DLight &lt = S.lights[__float_as_int(b.x)];
    const uint32_t m0 = B.amask[2u * g], m1 = B.amask[2u * g + 1u];
    B.amask[2u * g] = 0u;  // leave the mask clear for the next depth / wave
    B.amask[2u * g + 1u] = 0u;
    int spp = QB_AREA_SPP_MIN, s = 0;
    float inshadow = 0.0f;
    while (s < spp) {
      const V3 dir = wf_area_dir(lt, p, __float_as_uint(a.w), s);
      const float vis = (((s < 32 ? m0 : m1) >> (s & 31)) & 1u) ? 1.0f : 0.0f;
      const float falloff = P.no_falloff ? 1.f : qmin(1.f, 1.f / length2(dir));  // InverseSquareFalloff (lights.cpp:23-30)
      inshadow += (vis - inshadow) * falloff / (float) (s + 1);
      s++;
      if (inshadow > 0.f && inshadow < 1.f) spp = QB_AREA_SPP_MAX;
    }
    if (inshadow != 0.f) {
      float *acc = P.accum + 4ll * (long long) __float_as_int(b.y);
      atomicAdd(acc + 0, c.x * inshadow);
      atomicAdd(acc + 1, c.y * inshadow);
      atomicAdd(acc + 2, c.z * inshadow);
    }
  }
}

// ---- C-ABI ray-buffer kernels: the same traversal core over caller-supplied rays ----------------------------------------
__device__ __forceinline__ void unpack_ray(const qgRay &q, Ray3 &c, Ray3 &x, Ray3 &y)
{
  c.p = v3(q.p[0], q.p[1], q.p[2]);
  c.d = v3(q.dir[0], q.dir[1], q.dir[2]);
  if (q.has_diff) {
    x.p = v3(q.xp[0], q.xp[1], q.xp[2]);
    x.d = v3(q.xdir[0], q.xdir[1], q.xdir[2]);
    y.p = v3(q.yp[0], q.yp[1], q.yp[2]);
    y.d = v3(q.ydir[0], q.ydir[1], q.ydir[2]);
  } else {
    x = c;  // DiffRay(p, d): x and y are copies of the centre ray (src/core/ray.h:51-56)
    y = c;
  }
}

struct RayBufClosestIO {
  const qgRay *rays;
  uint32_t n;
  uint32_t *cur;
  float4 *r0, *r1;
  RedoQueue Q;
  __device__ uint32_t count() const { return n; }
  __device__ uint32_t *cursor() { return cur; }
  __device__ uint32_t map(uint32_t q) const { return q; }
  __device__ bool exact(uint32_t, int) const { return false; }  // true: walk this mesh with the degenerate-axis cull off
  __device__ bool defer(uint32_t idx, int k, float z_entry) const { return redo_push(Q, idx, k, z_entry); }
  __device__ bool load(uint32_t idx, Ray3 &c, float &tmax)
  {
    tmax = QB_BIGFLOAT;
    c.p = v3(rays[idx].p[0], rays[idx].p[1], rays[idx].p[2]);
    c.d = v3(rays[idx].dir[0], rays[idx].dir[1], rays[idx].dir[2]);
    return true;
  }
  __device__ void store(uint32_t idx, const HitRec &h) { put_rec(r0, r1, idx, h); }
  __device__ void store_any(uint32_t, bool) {}
  __device__ void store_redo(uint32_t idx, const HitRec &h) { put_rec_keep_w(r0, r1, idx, h); }
  __device__ void flag(uint32_t idx, int k) const { if (k < 32) atomicOr((uint32_t *) &r1[idx].w, 1u << k); }
  __device__ uint32_t flags(uint32_t idx) const { return __float_as_uint(r1[idx].w); }
};

struct RayBufAnyIO {
  const float *rays6, *tmaxs;
  uint32_t n;
  uint32_t *cur;
  uint8_t *occ;
  AnyQueue Q;
  __device__ AnyQueue any_queue() const { return Q; }
  __device__ uint32_t count() const { return n; }
  __device__ uint32_t *cursor() { return cur; }
  __device__ uint32_t map(uint32_t q) const { return q; }                 // queue position -> ray index
  __device__ bool exact(uint32_t, int) const { return false; }  // true: walk this mesh with the degenerate-axis cull off
  __device__ bool defer(uint32_t idx, int, float) const { return any_push(Q, idx); }
  __device__ bool load(uint32_t idx, Ray3 &c, float &tmax)
  {
    c.p = v3(rays6[6 * idx + 0], rays6[6 * idx + 1], rays6[6 * idx + 2]);
    c.d = v3(rays6[6 * idx + 3], rays6[6 * idx + 4], rays6[6 * idx + 5]);
    tmax = tmaxs[idx];
    return true;
  }
  __device__ void store(uint32_t, const HitRec &) {}
  __device__ void store_any(uint32_t idx, bool occluded) { occ[idx] = occluded ? 1 : 0; }
};

__global__ void QB_TRACE_BOUNDS k_trace_closest_rays(const DScene S, RayBufClosestIO io, unsigned long long *stats)
{
  TraceCounters cnt = {0, 0, 0};
  QB_TRAVERSE(false, io, cnt);
  flush_counters(stats, cnt);
}

__global__ void __launch_bounds__(QB_BLOCK) k_finish_rays(const DScene S, const qgRay *__restrict__ rays, uint32_t n, const float4 *r0,
                                                          const float4 *r1, qgHit *__restrict__ out)
{
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
    Ray3 c, x, y;
    unpack_ray(rays[idx], c, x, y);
    const HitRec rec = get_rec(r0, r1, idx);
    qgHit o;
    memset(&o, 0, sizeof o);
    o.hit = rec.k >= 0;
    o.node = -1;
    o.face = -1;
    o.z = rec.z;
    if (rec.k >= 0) {
      HitAttr h;
      // objects always see hasDiffRay = true (Node::ToNodeCoords returns a default-constructed DiffRay), so the
      // differential branch runs for every ray
      finish_hit(S, rec, c, x, y, true, true, h);
      o.node = S.obj_node[rec.k];
      o.face = h.face;
      o.mtl_id = h.mtl_id;
      o.front = h.front;
      o.has_texture = h.has_texture;
      o.p[0] = h.p.x; o.p[1] = h.p.y; o.p[2] = h.p.z;
      o.N[0] = h.N.x; o.N[1] = h.N.y; o.N[2] = h.N.z;
      o.uvw[0] = h.uvw.x; o.uvw[1] = h.uvw.y; o.uvw[2] = h.uvw.z;
      o.duvw[0] = h.duvw0.x; o.duvw[1] = h.duvw0.y; o.duvw[2] = h.duvw0.z;
      o.duvw[3] = h.duvw1.x; o.duvw[4] = h.duvw1.y; o.duvw[5] = h.duvw1.z;
    }
    out[idx] = o;
  }
}

__global__ void QB_TRACE_BOUNDS_ANY k_trace_any_rays(const DScene S, RayBufAnyIO io, unsigned long long *stats)
{
  TraceCounters cnt = {0, 0, 0};
  QB_TRAVERSE(true, io, cnt);
  flush_counters(stats, cnt, 5);
}

__global__ void __launch_bounds__(QB_ANYX_WARPS * 32) k_any_exact_rays(const DScene S, RayBufAnyIO io, unsigned long long *stats)
{
  __shared__ uint32_t s_stack[QB_ANYX_WARPS * QB_ANYX_STACK];
  if (*io.Q.count == 0u) return;
  TraceCounters cnt = {0, 0, 0};
  any_exact_pass(S, io, io.Q, s_stack, cnt);
  flush_counters(stats, cnt, 5);
}

__global__ void k_primary_rays(const WfParams P, size_t n, qgRay *__restrict__ out)
{
  for (size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t) gridDim.x * blockDim.x) {
    Ray3 c, x, y;
    primary_ray(P, P.pix_list[3 * idx], P.pix_list[3 * idx + 1], P.pix_list[3 * idx + 2], c, x, y);
    qgRay q;
    q.p[0] = c.p.x; q.p[1] = c.p.y; q.p[2] = c.p.z; q.dir[0] = c.d.x; q.dir[1] = c.d.y; q.dir[2] = c.d.z;
    q.xp[0] = x.p.x; q.xp[1] = x.p.y; q.xp[2] = x.p.z; q.xdir[0] = x.d.x; q.xdir[1] = x.d.y; q.xdir[2] = x.d.z;
    q.yp[0] = y.p.x; q.yp[1] = y.p.y; q.yp[2] = y.p.z; q.ydir[0] = y.d.x; q.ydir[1] = y.d.y; q.ydir[2] = y.d.z;
    q.has_diff = 1;
    out[idx] = q;
  }
}

// ---- launchers ------------------------------------------------------------------------------------------------------
static int persistent_grid(const void *kernel)
{
  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, QB_BLOCK, 0);
  if (per_sm < 1) per_sm = 1;
  return sms * per_sm;
}
#if QB_BLOCKPOOL
// Traversal kernels run over the block pool: QB_BP_THREADS threads and bp_smem_bytes(stack entries of the scene) of dynamic
// shared memory per block, two blocks per SM. The opt-in above 48 KB is per kernel and per device.
static size_t trace_smem(const DScene &S) { return bp_smem_bytes(S.bvh_stack); }
template <class K> static void trace_attr(K kernel)
{
  // keyed by (kernel, device): instantiations of one template share their function type, so a per-type flag would not do
  static std::mutex mu;
  static std::set<std::pair<const void *, int>> done;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!done.insert({(const void *) kernel, dev}).second) return;
  cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, QB_BP_SMEM);
}
static int trace_grid_for(size_t rays)  // persistent grid, smaller when the launch holds fewer rays than the pools
{
  const size_t full = (size_t) qb_sms() * QB_BP_BLOCKS;
  const size_t need = (rays + QB_BP_THREADS - 1) / QB_BP_THREADS;
  return (int) std::max<size_t>(1, std::min(full, need));
}
int wf_trace_grid() { return qb_sms() * QB_BP_BLOCKS; }
int wf_shadow_grid() { return qb_sms() * QB_BP_BLOCKS; }
#define QB_LAUNCH_TRACE(kernel, grid, st, ...)                                         \
  do {                                                                                 \
    trace_attr(kernel);                                                                \
    kernel<<<(grid), QB_BP_THREADS, trace_smem(S), (st)>>>(__VA_ARGS__);               \
  } while (0)
#else
static size_t trace_smem(const DScene &S) { return lanepool_smem_bytes(S.bvh_stack); }
// resident blocks of a traversal kernel for the scene's shared-memory need, cached per (kernel, bytes, device)
template <class K> static int trace_blocks_per_sm(K kernel, size_t smem)
{
  static std::mutex mu;
  static std::map<std::tuple<const void *, size_t, int>, int> cache;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(mu);
  const auto key = std::make_tuple((const void *) kernel, smem, dev);
  auto it = cache.find(key);
  if (it != cache.end()) return it->second;
  // the opt-in above 48 KB is one value per (kernel, device): only ever raised, so that a shallower scene's query does not
  // take away what a deeper scene's later launches (answered from the cache above) still need
  static std::map<std::pair<const void *, int>, size_t> opted;
  size_t &have = opted[std::make_pair((const void *) kernel, dev)];
  if (smem > 48 * 1024 && smem > have) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    have = smem;
  }
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, QB_BLOCK, smem);
  if (per_sm < 1) per_sm = 1;
  cache[key] = per_sm;
  return per_sm;
}
static int trace_grid_for(size_t rays)
{
  size_t g = (rays + QB_BLOCK - 1) / QB_BLOCK;
  const size_t cap = (size_t) qb_sms() * 8;
  return (int) std::max<size_t>(1, std::min(g, cap));
}
// does a stack of `entries + extra` entries keep as many blocks of the closest-hit kernel resident as one of `entries`?
bool trace_stack_fits(int entries, int extra)
{
  return trace_blocks_per_sm(k_wf_trace<false>, lanepool_smem_bytes(entries + extra)) == trace_blocks_per_sm(k_wf_trace<false>, lanepool_smem_bytes(entries));
}
int wf_trace_grid() { return 0; }   // persistent grids are sized at launch: they depend on the scene's stack depth
int wf_shadow_grid() { return 0; }
// grid <= 0: persistent grid = SMs x resident blocks
#define QB_LAUNCH_TRACE(kernel, grid, st, ...)                                                        \
  do {                                                                                                \
    const size_t smem_ = trace_smem(S);                                                               \
    const int per_sm_ = trace_blocks_per_sm(kernel, smem_);                                           \
    const int grid_ = (grid) > 0 ? std::min((int) (grid), qb_sms() * per_sm_) : qb_sms() * per_sm_;   \
    kernel<<<grid_, QB_BLOCK, smem_, (st)>>>(__VA_ARGS__);                                            \
  } while (0)
#endif

void launch_wf_trace(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st)
{
  // traversal by the ray pool, then the hit attributes by full warps
  // grid-stride kernels sized to what is resident at once: a larger grid runs its last blocks alone at low occupancy.
  // k_wf_finish is a chain of dependent loads (record -> object -> mesh -> face -> normals / uv): capped at 64 registers
  // (QB_FINISH_BLOCKS 8 blocks / SM, 112-136 B of spills) it hides more of that latency than at its natural 96 registers
  // and 5 blocks — A/B on the tower: 73.96 -> 72.14 ms per step (6 blocks: 72.34)
  static const int fg_primary = persistent_grid((const void *) k_wf_finish<true>), fg_secondary = persistent_grid((const void *) k_wf_finish<false>);
  const int fg = P.depth == 0 ? fg_primary : fg_secondary;
  const int redo_grid = QB_BLOCKPOOL ? qb_sms() / 4 : qb_sms();
  if (P.depth == 0) {
    QB_LAUNCH_TRACE(k_wf_trace<true>, grid, st, S, B, P);
    k_redo_detect<true><<<qb_sms() * 8, QB_BLOCK, 0, st>>>(S, B, P);
    QB_LAUNCH_TRACE(k_redo_exact<true>, redo_grid, st, S, B, P);
    if (!P.fused_finish) k_wf_finish<true><<<fg, QB_BLOCK, 0, st>>>(S, B, P);
  } else {
    QB_LAUNCH_TRACE(k_wf_trace<false>, grid, st, S, B, P);
    k_redo_detect<false><<<qb_sms() * 8, QB_BLOCK, 0, st>>>(S, B, P);
    QB_LAUNCH_TRACE(k_redo_exact<false>, redo_grid, st, S, B, P);
    if (!P.fused_finish) k_wf_finish<false><<<fg, QB_BLOCK, 0, st>>>(S, B, P);
  }
}
void launch_wf_shadow(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st)
{
  QB_LAUNCH_TRACE(k_wf_shadow, grid, st, S, B, P);
  k_any_exact_shadow<<<qb_sms() * 4, QB_ANYX_WARPS * 32, 0, st>>>(S, B, P);
}
void launch_wf_area(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st)
{
  QB_LAUNCH_TRACE(k_wf_area_shadow, grid, st, S, B, P);
  k_any_exact_area<<<qb_sms() * 4, QB_ANYX_WARPS * 32, 0, st>>>(S, B, P);
  k_wf_area_resolve<<<qb_sms() * 8, QB_BLOCK, 0, st>>>(S, B, P);
}
static int flat_grid(size_t n)
{
  size_t g = (n + QB_BLOCK - 1) / QB_BLOCK;
  const size_t cap = (size_t) qb_sms() * 8;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int) g;
}
__global__ void __launch_bounds__(QB_BLOCK) k_redo_detect_rays(const DScene S, RayBufClosestIO io)
{
  if (*io.Q.count == 0u) return;
  redo_detect(S, io, io.Q);
}
__global__ void QB_TRACE_BOUNDS k_redo_exact_rays(const DScene S, RayBufClosestIO base, unsigned long long *stats)
{
  const uint32_t n = min(*base.Q.count, base.Q.cap);
  if (n == 0u) return;
  TraceCounters cnt = {0, 0, 0};
  RedoIO<RayBufClosestIO> io{base, base.Q, n};
  QB_TRAVERSE(false, io, cnt);
  flush_counters(stats, cnt);
}
// scratch: a 16-byte header (work cursor, cull-check count, cull-check cursor), two float4 records per ray, then the
// cull-check queue (one uint4 per ray); the header is zeroed by the launch
size_t trace_rays_scratch_bytes(size_t n) { return 16 + 48 * n; }
void launch_trace_closest_rays(const DScene &S, const qgRay *rays, size_t n, qgHit *out, unsigned long long *stats, void *scratch,
                               cudaStream_t st)
{
  uint32_t *cur = (uint32_t *) scratch;
  float4 *r0 = (float4 *) ((char *) scratch + 16), *r1 = r0 + n;
  cudaMemsetAsync(cur, 0, 16, st);
  // QB_REDO_CAP: test hook — a smaller cull-check queue than one entry per ray, so the overflow path (the lane repeats the
  // mesh itself, first pass and redo pass alike) is exercised by the edge fixtures (tests/test_gpu_edges.py)
  uint32_t redo_cap = (uint32_t) n;
  if (const char *e = getenv("QB_REDO_CAP")) redo_cap = (uint32_t) std::min<size_t>(n, (size_t) std::max(0, atoi(e)));
  RayBufClosestIO io{rays, (uint32_t) n, cur, r0, r1, RedoQueue{(uint4 *) (r1 + n), redo_cap, cur + 1, cur + 2}};
  QB_LAUNCH_TRACE(k_trace_closest_rays, trace_grid_for(n), st, S, io, stats);
  k_redo_detect_rays<<<qb_sms() * 8, QB_BLOCK, 0, st>>>(S, io);
  QB_LAUNCH_TRACE(k_redo_exact_rays, QB_BLOCKPOOL ? qb_sms() / 4 : qb_sms(), st, S, io, stats);
  k_finish_rays<<<flat_grid(n), QB_BLOCK, 0, st>>>(S, rays, (uint32_t) n, r0, r1, out);
}
void launch_trace_any_rays(const DScene &S, const float *rays6, const float *tmax, size_t n, uint8_t *occ, unsigned long long *stats,
                           void *scratch, cudaStream_t st)
{
  uint32_t *cur = (uint32_t *) scratch;
  cudaMemsetAsync(cur, 0, 16, st);
  // scratch: 16-byte header (work cursor, hand-over count, hand-over cursor), then the hand-over queue (one index per ray)
  // QB_ANYQ_CAP: test hook — a smaller queue, so the in-lane repeat of a full queue is exercised as well
  uint32_t qcap = (uint32_t) n;
  if (const char *e = getenv("QB_ANYQ_CAP")) qcap = (uint32_t) std::min<size_t>(n, (size_t) std::max(0, atoi(e)));
  RayBufAnyIO io{rays6, tmax, (uint32_t) n, cur, occ, AnyQueue{(uint32_t *) ((char *) scratch + 16), qcap, cur + 1, cur + 2}};
  QB_LAUNCH_TRACE(k_trace_any_rays, trace_grid_for(n), st, S, io, stats);
  k_any_exact_rays<<<qb_sms() * 4, QB_ANYX_WARPS * 32, 0, st>>>(S, io, stats);
}
void launch_primary_rays(const WfParams &P, size_t n, qgRay *out, cudaStream_t st)
{
  k_primary_rays<<<flat_grid(n), QB_BLOCK, 0, st>>>(P, n, out);
}

// ---- any-hit batch without an allocation (qg_trace_any_noalloc) --------------------------------------------------------------
// The scratch of launch_trace_any_rays (16-byte header + one queue entry per ray) lives in a file-scope device array instead of
// a cudaMalloc'ed buffer: header in the first 256 bytes, the hand-over queue of the warp-per-ray exact walk after it. Every
// GPU that loads the library holds the 4 MiB + 256 B until exit. All calls on a device share it: calls that overlap on the GPU
// (two streams, two threads) race on the cursor and the queue.
static __device__ __align__(256) unsigned char g_noalloc_scratch[256 + 4 * (size_t) QB_NOALLOC_MAX_RAYS];

// The array's name is a host-side stand-in in host code, not a device address: ask the runtime for the address once per device,
// with that device current (the caller has made it current: use_scene).
static cudaError_t noalloc_scratch(unsigned char **out)
{
  static std::mutex mu;
  static unsigned char *addr[QB_NOALLOC_MAX_DEVICES] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= QB_NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void *p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    addr[dev] = (unsigned char *) p;
  }
  *out = addr[dev];
  return cudaSuccess;
}

cudaError_t launch_trace_any_rays_noalloc(const DScene &S, const float *rays6, const float *tmax, size_t n, uint8_t *occ,
                                          unsigned long long *stats, cudaStream_t st)
{
  if (n > (size_t) QB_NOALLOC_MAX_RAYS) return cudaErrorInvalidValue;
  unsigned char *scratch = nullptr;
  cudaError_t e = noalloc_scratch(&scratch);
  if (e != cudaSuccess) return e;
  uint32_t *cur = (uint32_t *) scratch;
  // the array keeps the last call's contents: the work cursor, the hand-over count and its cursor must start at zero; the
  // queue entries below the count are all written before they are read
  e = cudaMemsetAsync(cur, 0, 16, st);
  if (e != cudaSuccess) return e;
  RayBufAnyIO io{rays6, tmax, (uint32_t) n, cur, occ, AnyQueue{(uint32_t *) (scratch + 256), (uint32_t) n, cur + 1, cur + 2}};
  QB_LAUNCH_TRACE(k_trace_any_rays, trace_grid_for(n), st, S, io, stats);
  k_any_exact_rays<<<qb_sms() * 4, QB_ANYX_WARPS * 32, 0, st>>>(S, io, stats);
  return cudaGetLastError();
}

}  // namespace qb
