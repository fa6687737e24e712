// wavefront.h — buffers and parameters of the wavefront pipeline that replaces the reference's TBB/OpenMP tile
// loop (src/tasking/parallel_for.cpp:75-86 driving Renderer::ThreadRender, src/renderers/renderer.cpp:370-423).
//
// One WAVE = up to `capacity` samples (pixel, sample-index pairs) in flight. Per wave, for depth d = 0..maxBounce:
//     trace_closest(d)  ->  shade(d)  ->  trace_shadow(d)
// trace_closest(0) generates the primary rays itself (ray generation fused, no ray buffer read).
// shade(d) compacts surviving paths into the state buffer of depth d+1 and appends next-event shadow rays to the
// shadow queue with warp-aggregated atomics. All kernels are persistent: a fixed grid of warps pulls 32-ray
// batches from a device-side work counter, so no host round trip is needed between bounces.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "dev_scene.h"

namespace qb {

#define QB_MAX_BOUNCE 15  // depth slots in the counter block

// path flags (low 8 bits of ps2.w)
#define QB_FLAG_DIFFUSE_HIT 1u  // HitInfo::hasDiffuseHit of MtlBlinn_PhotonMap
#define QB_FLAG_HAS_DIFF 2u     // DiffRay::hasDiffRay: ps4..ps6 hold this path's differential rays

// device counter block (uint32 each), zeroed at wave start
//   [0 .. 15]   n_paths[d]    paths queued for depth d (d >= 1; depth 0 uses the wave size)
//   [16 .. 31]  n_shadow[d]   shadow rays queued by shade(d)
//   [32 .. 47]  fetch_trace[d], [48 .. 63] fetch_shade[d], [64 .. 79] fetch_shadow[d]   work-fetch cursors
#define QB_CNT_PATHS 0
#define QB_CNT_SHADOW 16
#define QB_CNT_FETCH_TRACE 32
#define QB_CNT_FETCH_SHADE 48
#define QB_CNT_FETCH_SHADOW 64
#define QB_CNT_WORDS 96

struct WfBuffers {
  // path state, ping-pong by depth parity: [d & 1]
  float4 *ps0[2];  // origin.xyz, dir.x
  float4 *ps1[2];  // dir.yz, accum slot (int bits), rng key (uint bits)
  float4 *ps2[2];  // throughput.rgb, packed (rng counter << 8 | flags)
  float4 *ps3[2];  // absorption.rgb of the medium behind the spawning surface, unused
  // differential rays of secondary paths — only the Whitted-style material classes carry them (MtlBlinn_Basic.cpp:131,
  // MtlBlinn_MonteCarloGI.cpp:132); null for the path-tracing classes, whose secondary rays are DiffRay(p, dir)
  float4 *ps4[2];  // x.p.xyz, x.dir.x
  float4 *ps5[2];  // x.dir.yz, y.p.xy
  float4 *ps6[2];  // y.p.z, y.dir.xyz
  // hit records written by trace_closest(d), read by shade(d)
  float4 *h0;  // t, object node (int, -1 miss), mtl_id (int), flags (front | has_texture << 1)
  float4 *h1;  // p.xyz, N.x
  float4 *h2;  // N.yz, u, v
  float4 *h3;  // du0, dv0, du1, dv1
  float4 *h4;  // DiffHitInfo x.z, y.z (hit distances of the differential rays); allocated with ps4..ps6
  // shadow queue
  float4 *s0;  // origin.xyz, dir.x
  float4 *s1;  // dir.yz, tmax, accum slot (int bits)
  float4 *s2;  // contribution.rgb if unoccluded
  uint32_t *counters;
  unsigned long long *stats;  // [0] closest rays, [1] shadow rays, [2] nodes, [3] tris, [4] objects
  uint32_t capacity;          // paths per wave
  uint32_t shadow_capacity;   // capacity * number of non-ambient lights
};

struct WfParams {
  // camera frame (Renderer::ComputeScene, renderer.cpp:77-91)
  V3 cam_pos, scr_u, scr_v, scr_a;
  int32_t width, height;
  // sample enumeration
  int32_t mode;              // 0: 4x4 tiles k = tile_rank + lt*tile_ranks (renderer.cpp:383-398); 1: explicit (i,j,s) list
  int32_t tiles_x, tile_rank, tile_ranks, tiles_local;
  int32_t sample_begin;
  unsigned long long wave_base;  // index of the wave's first path in the shard's enumeration
  const int32_t *pix_list;   // mode 1: i,j,s triples
  uint32_t n_paths;          // paths in this wave
  uint32_t seed;
  int32_t max_bounce;
  int32_t depth;             // current depth d
  int32_t stale_tex;
  int32_t mtl_class;         // QG_CLASS_*: which of the reference's MtlBlinn classes shades (materials.h:57-61)
  int32_t carry_diff;        // secondary paths carry differential rays (ps4..ps6 / h4 allocated)
  int32_t no_falloff;        // DisableInverseSquareFalloff() was called (MtlBlinn_Basic.cpp:18, lights.cpp:21-30)
  float *accum;              // float4 per slot: sum r,g,b, weight
  float *depth_plane;        // may be null
};

// decode path index -> pixel (i,j), sample s, accumulation slot; false for padding lanes of edge tiles
__host__ __device__ inline bool wf_decode(const WfParams &P, uint32_t idx, int &i, int &j, int &s, int &slot)
{
  if (P.mode == 1) {
    i = P.pix_list[3 * idx + 0];
    j = P.pix_list[3 * idx + 1];
    s = P.pix_list[3 * idx + 2];
    slot = (int) idx;
    return true;
  }
  const unsigned long long g = P.wave_base + idx;
  const unsigned long long per_s = (unsigned long long) P.tiles_local * 16ull;
  s = P.sample_begin + (int) (g / per_s);
  const uint32_t rem = (uint32_t) (g % per_s);
  const uint32_t lt = rem >> 4, pi = rem & 15u;
  const uint32_t k = (uint32_t) P.tile_rank + lt * (uint32_t) P.tile_ranks;
  const uint32_t tx = k % (uint32_t) P.tiles_x, ty = k / (uint32_t) P.tiles_x;
  i = (int) (tx * 4u + (pi & 3u));
  j = (int) (ty * 4u + (pi >> 2));
  slot = j * P.width + i;
  return i < P.width && j < P.height;
}

// launch wrappers (implemented in trace_kernels.cu / shade_kernels.cu)
void launch_wf_trace(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st);
void launch_wf_shadow(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st);
void launch_wf_shade(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st);
void launch_wf_stats(const WfBuffers &B, int max_bounce, cudaStream_t st);
void launch_resolve(const float *accum, const float *depth_in, int n, int spp, int srgb, float *rgb_mean, uint8_t *rgb8, float *depth_out,
                    uint8_t *sample_count, uint8_t *mask, cudaStream_t st);
size_t trace_rays_scratch_bytes(size_t n);  // device scratch the two ray-buffer launchers need
void launch_trace_closest_rays(const DScene &S, const qgRay *rays, size_t n, qgHit *out, unsigned long long *stats, void *scratch,
                               cudaStream_t st);
void launch_trace_any_rays(const DScene &S, const float *rays6, const float *tmax, size_t n, uint8_t *occ, unsigned long long *stats,
                           void *scratch, cudaStream_t st);
void launch_primary_rays(const WfParams &P, size_t n, qgRay *out, cudaStream_t st);
int wf_trace_grid();   // persistent grid sizes (blocks), from the occupancy calculator
int wf_shadow_grid();
int wf_shade_grid();
void upload_texture_taps(const float *xy64);  // 32 (x,y) tap offsets of Texture::Sample's elliptical filter


}  // namespace qb
