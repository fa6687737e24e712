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
#include "trace.cuh"
#include "wavefront.h"
#include <algorithm>

namespace qb {

__device__ __forceinline__ float4 f4(float x, float y, float z, float w) { return make_float4(x, y, z, w); }

// primary DiffRay of pixel (i,j), sample s (no depth of field): texpos = (i + H11(s), j + H13(s))
__device__ __forceinline__ void primary_ray(const WfParams &P, int i, int j, int s, Ray3 &c, Ray3 &x, Ray3 &y)
{
  const float tx = halton(s, 11) + (float) i;
  const float ty = halton(s, 13) + (float) j;
  const V3 cpt = P.scr_a + tx * P.scr_u + ty * P.scr_v;
  const V3 xpt = P.scr_a + (tx + 0.01f) * P.scr_u + ty * P.scr_v;
  const V3 ypt = P.scr_a + tx * P.scr_u + (ty + 0.01f) * P.scr_v;
  c.p = x.p = y.p = P.cam_pos;
  c.d = normalize(cpt - P.cam_pos);
  x.d = normalize(xpt - P.cam_pos);
  y.d = normalize(ypt - P.cam_pos);
}

// counting build only: stats[2..4] = BVH nodes / triangles / objects visited by closest-hit rays, [5..7] by any-hit rays
__device__ __forceinline__ void flush_counters(unsigned long long *stats, const TraceCounters &cnt, int base = 2)
{
#ifdef QB_COUNT
  atomicAdd(&stats[base + 0], cnt.nodes);
  atomicAdd(&stats[base + 1], cnt.tris);
  atomicAdd(&stats[base + 2], cnt.objects);
#endif
}

// PRIMARY = depth 0 (rays generated in registers) / depth >= 1 (rays read from the path state). Two instantiations
// rather than a run-time branch: with both ray sources in one kernel nvcc 12.9 merged the two definitions of the
// origin's x component into the camera's (the state load of .x was dropped), found by the radiance parity test.
template <bool PRIMARY>
__global__ void __launch_bounds__(QB_BLOCK) k_wf_trace(const DScene S, const WfBuffers B, const WfParams P)
{
  const int d = P.depth;
  const uint32_t n = PRIMARY ? P.n_paths : B.counters[QB_CNT_PATHS + d];
  const int lane = threadIdx.x & 31;
  const int pb = d & 1;
  TraceCounters cnt = {0, 0, 0};
  for (;;) {
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(&B.counters[QB_CNT_FETCH_TRACE + d], 32u);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base >= n) break;
    const uint32_t idx = base + lane;
    if (idx >= n) continue;
    Ray3 c, x, y;
    const bool with_diff = PRIMARY;
    if (PRIMARY) {
      int i, j, s, slot;
      if (!wf_decode(P, idx, i, j, s, slot)) {
        B.h0[idx] = f4(QB_BIGFLOAT, __int_as_float(-2), 0.f, 0.f);  // padding lane of an edge tile
        continue;
      }
      primary_ray(P, i, j, s, c, x, y);
    } else {
      const float4 a = B.ps0[pb][idx], b = B.ps1[pb][idx];
      c.p = v3(a.x, a.y, a.z);
      c.d = v3(a.w, b.x, b.y);
      x = c;
      y = c;
    }
    float z;
    HitAttr h;
    const int k = trace_closest(S, c, x, y, with_diff, P.stale_tex != 0, z, h, cnt);
    if (k < 0) {
      B.h0[idx] = f4(QB_BIGFLOAT, __int_as_float(-1), 0.f, 0.f);
    } else {
      const int flags = (h.front ? 1 : 0) | (h.has_texture ? 2 : 0);
      B.h0[idx] = f4(z, __int_as_float(S.obj_node[k]), __int_as_float(h.mtl_id), __int_as_float(flags));
      B.h1[idx] = f4(h.p.x, h.p.y, h.p.z, h.N.x);
      B.h2[idx] = f4(h.N.y, h.N.z, h.uvw.x, h.uvw.y);
      B.h3[idx] = f4(h.duvw0.x, h.duvw0.y, h.duvw1.x, h.duvw1.y);
    }
  }
  flush_counters(B.stats, cnt);
}

__global__ void __launch_bounds__(QB_BLOCK) k_wf_shadow(const DScene S, const WfBuffers B, const WfParams P)
{
  const int d = P.depth;
  const uint32_t n = B.counters[QB_CNT_SHADOW + d];
  const int lane = threadIdx.x & 31;
  TraceCounters cnt = {0, 0, 0};
  for (;;) {
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(&B.counters[QB_CNT_FETCH_SHADOW + d], 32u);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (base >= n) break;
    const uint32_t idx = base + lane;
    if (idx >= n) continue;
    const float4 a = B.s0[idx], b = B.s1[idx];
    Ray3 r;
    r.p = v3(a.x, a.y, a.z);
    r.d = v3(a.w, b.x, b.y);
    if (!trace_any(S, r, b.z, cnt)) {
      const float4 c = B.s2[idx];
      float *acc = P.accum + 4ll * (long long) __float_as_int(b.w);
      atomicAdd(acc + 0, c.x);
      atomicAdd(acc + 1, c.y);
      atomicAdd(acc + 2, c.z);
    }
  }
  flush_counters(B.stats, cnt, 5);
}

// ---- C-ABI ray-buffer kernels -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(QB_BLOCK) k_trace_closest_rays(const DScene S, const qgRay *__restrict__ rays, size_t n,
                                                                 qgHit *__restrict__ out, unsigned long long *stats)
{
  TraceCounters cnt = {0, 0, 0};
  for (size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t) gridDim.x * blockDim.x) {
    const qgRay q = rays[idx];
    Ray3 c, x, y;
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
    float z;
    HitAttr h;
    // objects always see hasDiffRay = true (Node::ToNodeCoords returns a default-constructed DiffRay), so the
    // differential branch runs for every ray
    const int k = trace_closest(S, c, x, y, true, true, z, h, cnt);
    qgHit o;
    memset(&o, 0, sizeof o);
    o.hit = k >= 0;
    o.node = -1;
    o.face = -1;
    o.z = z;
    if (k >= 0) {
      o.node = S.obj_node[k];
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
  flush_counters(stats, cnt);
}

__global__ void __launch_bounds__(QB_BLOCK) k_trace_any_rays(const DScene S, const float *__restrict__ rays6,
                                                             const float *__restrict__ tmax, size_t n, uint8_t *__restrict__ occ,
                                                             unsigned long long *stats)
{
  TraceCounters cnt = {0, 0, 0};
  for (size_t idx = (size_t) blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (size_t) gridDim.x * blockDim.x) {
    Ray3 r;
    r.p = v3(rays6[6 * idx + 0], rays6[6 * idx + 1], rays6[6 * idx + 2]);
    r.d = v3(rays6[6 * idx + 3], rays6[6 * idx + 4], rays6[6 * idx + 5]);
    occ[idx] = trace_any(S, r, tmax[idx], cnt) ? 1 : 0;
  }
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
int wf_trace_grid()
{
  static int g = std::min(persistent_grid((const void *) k_wf_trace<true>), persistent_grid((const void *) k_wf_trace<false>));
  return g;
}
int wf_shadow_grid() { static int g = persistent_grid((const void *) k_wf_shadow); return g; }

void launch_wf_trace(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st)
{
  if (P.depth == 0) k_wf_trace<true><<<grid, QB_BLOCK, 0, st>>>(S, B, P);
  else k_wf_trace<false><<<grid, QB_BLOCK, 0, st>>>(S, B, P);
}
void launch_wf_shadow(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st)
{
  k_wf_shadow<<<grid, QB_BLOCK, 0, st>>>(S, B, P);
}
static int flat_grid(size_t n)
{
  size_t g = (n + QB_BLOCK - 1) / QB_BLOCK;
  if (g > 148 * 16) g = 148 * 16;
  if (g < 1) g = 1;
  return (int) g;
}
void launch_trace_closest_rays(const DScene &S, const qgRay *rays, size_t n, qgHit *out, unsigned long long *stats, cudaStream_t st)
{
  k_trace_closest_rays<<<flat_grid(n), QB_BLOCK, 0, st>>>(S, rays, n, out, stats);
}
void launch_trace_any_rays(const DScene &S, const float *rays6, const float *tmax, size_t n, uint8_t *occ, unsigned long long *stats,
                           cudaStream_t st)
{
  k_trace_any_rays<<<flat_grid(n), QB_BLOCK, 0, st>>>(S, rays6, tmax, n, occ, stats);
}
void launch_primary_rays(const WfParams &P, size_t n, qgRay *out, cudaStream_t st)
{
  k_primary_rays<<<flat_grid(n), QB_BLOCK, 0, st>>>(P, n, out);
}

}  // namespace qb
