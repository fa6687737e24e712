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
#define QB_CNT_AREA 80         // [80 .. 95] n_area[d]: (hit, area light) groups queued by shade(d)
#define QB_CNT_FETCH_AREA 96   // [96 .. 111] work-fetch cursor of the area-light shadow kernel
#define QB_CNT_REDO 112        // [112 .. 127] closest-hit rays of trace(d) whose culled mesh walk must be checked (k_redo_detect)
#define QB_CNT_FETCH_REDO 128  // [128 .. 143] work-fetch cursor of k_redo_exact
#define QB_CNT_HOLES_PATHS 144  // [144 .. 159] dead entries in the path queue of depth d (unused tails of the shade warps' reservations)
#define QB_CNT_HOLES_SHADOW 160 // [160 .. 175] dead entries in the shadow queue of depth d
#define QB_CNT_GATHER 176      // [176 .. 191] photon-map gather requests queued by shade(d) (-use-photon-map)
#define QB_CNT_ANYQ 192            // [192 .. 207] shadow rays of depth d handed to k_any_exact (anyexact.cuh)
#define QB_CNT_FETCH_ANYQ 208      // [208 .. 223] its work cursor
#define QB_CNT_ANYQ_AREA 224       // [224 .. 239] the same for the area-light shadow samples
#define QB_CNT_FETCH_ANYQ_AREA 240 // [240 .. 255]
#define QB_CNT_WORDS 256

// area lights (PointLight / SpotLight with size > 0.01, src/lights/lights.cpp:50-65,84-101): shadow samples per group
#ifndef QB_SHADE_CHUNK
#define QB_SHADE_CHUNK 256       // paths a shade warp reserves per cursor atomic (multiple of 32)
#endif
#ifndef QB_OUT_CHUNK
#define QB_OUT_CHUNK 256         // queue slots a shade warp reserves per push atomic (k_wf_shade); multiple of 32
#endif
#define QB_MAT_BINS 256        // material sort: bin 0 = miss, 1 = no material / MultiMtl out of range, 2 + i = material i
#define QB_AREA_SPP_MIN 16     // GenLight::shadow_spp_min (lights.cpp:16)
#define QB_AREA_SPP_MAX 64     // GenLight::shadow_spp_max (lights.cpp:17)

// Photon map as the gather kernel reads it (photon.cuh): cyPhotonMap's own left-balanced kd-tree (src/ext/cyPhotonMap.h:226-318),
// photon i of the tree at index i (index 0 unused, children of i at 2 i and 2 i + 1), dequantised the way
// Photon::GetPower / GetDirection return them (:61-70,186-206). The tree — not just any nearest-neighbour structure — is kept
// because LocatePhotons' result depends on the ORDER candidates arrive in (see photon.cuh).
struct DPhotonMap {
  const float4 *pa;   // position.xyz, max power (Photon::power after ScalePhotonPowers)
  const float4 *pb;   // direction.xyz, bits: colour r | g << 8 | b << 16, splitting plane << 24
  uint32_t count;     // photons (tree indices 1 .. count)
  int32_t half;       // halfStoredPhotons: nodes with index < half are descended (cyPhotonMap.h:247)
  float radius;       // PhotonMap::radius of the scene (renderer.cpp:123-127)
};

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
  // area-light groups: one entry per (shaded hit, area light); its 64 shadow rays are generated on the fly
  float4 *a0;       // p.xyz, group rng key (uint bits)
  float4 *a1;       // light index (int bits), accum slot (int bits), unused
  float4 *a2;       // contribution.rgb per unit of the light's visibility estimate
  uint32_t *amask;  // 2 words per group: bit s = shadow sample s reached the light
  uint32_t area_capacity;
  // deferred checks of the degenerate-axis cull (traverse.cuh): {ray index, object, z at mesh entry (bits), flag}
  uint4 *redo;
  uint32_t redo_capacity;
  // photon-map gather requests of the depth being shaded (allocated by the first render with use_photon_map)
  float4 *g0;  // p.xyz, N.x
  float4 *g1;  // N.yz, V.xy
  float4 *g2;  // V.z, diffuse.rgb
  float4 *g3;  // specular.rgb, glossiness
  float4 *g4;  // throughput.rgb, accum slot | maps << 30 (bit 30 photon map, bit 31 caustics map)
  // any-hit rays handed to the warp-per-ray exact walk (anyexact.cuh): ray indices
  uint32_t *anyq;
  uint32_t anyq_capacity;
  // material sort of the shade queue (optional): permutation and the per-material bins / offsets
  uint32_t *perm;      // capacity entries
  uint32_t *mat_bins;  // 2 * QB_MAT_BINS: counts, then running offsets
  uint32_t *counters;
  unsigned long long *stats;  // [0] closest rays, [1] shadow rays, [2] nodes, [3] tris, [4] objects
  uint32_t capacity;          // paths per wave
  uint32_t shadow_capacity;   // capacity * number of non-ambient lights
};

// A batch of emissions [e0, e0 + E). rays / path_of are compact ping-pong arrays of the photons still flying; everything else
// is indexed by the emission's position in the batch.
struct PhotonBatch {
  qgRay *rays[2];
  uint32_t *path_of[2];
  qgHit *hits;          // closest hits of rays[cur] (launch_trace_closest_rays)
  float4 *state;        // photon power r g b, packed: bounce | hasDiffuseHit << 8
  uint32_t *rng_n;      // draws consumed from the emission's stream
  uint32_t *count;      // photons recorded by the emission
  float *rec;           // [emission][max_bounce][9]: position, incoming direction, power
  uint32_t *n_next;     // survivors appended to rays[cur ^ 1]
  int32_t max_bounce;
};

struct WfParams {
  // camera frame (Renderer::ComputeScene, renderer.cpp:77-91)
  V3 cam_pos, scr_u, scr_v, scr_a;
  int32_t width, height;
  // depth of field (renderer.cpp:319-322): lens radius (active above 0.1) and the camera's X / Y axes
  float dof;
  V3 scr_x, scr_y;
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
  int32_t sort_mat;          // 1: shade(d) walks its queue in material order (B.perm)
  int32_t mtl_class;         // QG_CLASS_*: which of the reference's MtlBlinn classes shades (materials.h:57-61)
  int32_t carry_diff;        // secondary paths carry differential rays (ps4..ps6 / h4 allocated)
  int32_t no_falloff;        // DisableInverseSquareFalloff() was called (MtlBlinn_Basic.cpp:18, lights.cpp:21-30)
  int32_t fused_finish;      // 1: no k_wf_finish launch, k_wf_shade<true> computes the hit attributes from trace's record itself
  int32_t use_photon_map;    // Scene::usePhotonMap: diffuse hits gather from pm[0] (photon map) / pm[1] (caustics map)
  DPhotonMap pm[2];
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

// Ray origin of pixel (i,j), sample s: the camera position, or with depth of field a point on the lens disc drawn with
// the sample's first two random numbers (SuperSamplerHalton::NewDofSample, src/scene/scene.cpp:104-111).
// wf_dof_draws = how many numbers of the sample's stream that consumed.
__host__ __device__ inline uint32_t wf_dof_draws(const WfParams &P) { return P.dof > 0.1f ? 2u : 0u; }
__host__ __device__ inline V3 wf_campos(const WfParams &P, int i, int j, int s)
{
  if (!(P.dof > 0.1f)) return P.cam_pos;
  const uint32_t key = rng_key((uint32_t) (j * P.width + i), (uint32_t) s, P.seed);
  const float r1 = rng_float(key, 0u), r2 = rng_float(key, 1u);
  const float r = P.dof * sqrtf(r1);
  const float t = r2 * 2.f * 3.14159274101257324f;
  return P.cam_pos + ((r * cosf(t)) * P.scr_x + (r * sinf(t)) * P.scr_y);
}

// Shadow sample s of an area-light group: the un-normalised direction `position + UniformBall(size) - p` of
// PointLight::Illuminate (lights.cpp:54). Every sample has its own counter-based stream keyed by (group key, s), so the
// shadow kernel and the resolve kernel regenerate the same direction without storing 64 rays per group.
// Sampler::UniformBall (core/sampler.cpp:42-53): rejection from the cube, z re-uses r2 as the reference does.
__host__ __device__ inline V3 wf_area_dir(const DLight &lt, V3 p, uint32_t gkey, int s)
{
  const uint32_t key = rng_mix(gkey ^ rng_mix((uint32_t) s * 0x9E3779B9u + 0x51ED270Bu));
  uint32_t n = 0;
  V3 b = v3(0.f);
  for (int it = 0; it < 64; ++it) {
    const float r1 = rng_float(key, n), r2 = rng_float(key, n + 1u);
    n += 3u;  // Get3f draws three numbers; the third is unused (sampler.cpp:50)
    b = v3((2.f * r1 - 1.f) * lt.size, (2.f * r2 - 1.f) * lt.size, (2.f * r2 - 1.f) * lt.size);
    if (!(length(b) > lt.size)) break;
  }
  return (v3(lt.position[0], lt.position[1], lt.position[2]) + b) - p;
}

// launch wrappers (implemented in trace_kernels.cu / shade_kernels.cu)
void launch_wf_trace(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st);
void launch_wf_shadow(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st);
void launch_wf_shade(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st);
void launch_wf_area(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st);  // area shadow rays + resolve
void launch_mat_sort(const DScene &S, const WfBuffers &B, const WfParams &P, cudaStream_t st);  // builds B.perm for shade(P.depth)
void launch_wf_stats(const WfBuffers &B, int max_bounce, cudaStream_t st);
void launch_photon_gather(const DScene &S, const WfBuffers &B, const WfParams &P, cudaStream_t st);  // the gather requests of shade(P.depth)
void launch_photon_emit(const DScene &S, const PhotonBatch &Bt, uint32_t e0, uint32_t E, int map, uint32_t seed, int n_plights, const int32_t *plights,
                        cudaStream_t st);
void launch_photon_step(const DScene &S, const PhotonBatch &Bt, int cur, uint32_t n_cur, int map, uint32_t e0, uint32_t seed, cudaStream_t st);
void launch_photon_estimate(const DPhotonMap &M, const float *pos, const float *nrm, uint32_t n, float *irrad, float *dir, cudaStream_t st);
// adaptive sampling (SuperSamplerHalton, src/scene/scene.cpp:83-123), see shade_kernels.cu
void launch_adaptive_build(int W, int H, int tile_rank, int tile_ranks, const int32_t *count, const uint8_t *active, int chunk,
                           uint32_t *n_active, int32_t *list, int32_t *triples, cudaStream_t st);
void launch_adaptive_update(uint32_t n_active, const int32_t *list, int chunk, const float *rad, const float *rad_depth, int spp_min,
                            int spp_max, int32_t *count, uint8_t *active, float *mean, float *sdev, float *depth, cudaStream_t st);
void launch_adaptive_finish(int n, const int32_t *count, const float *mean, float *accum, cudaStream_t st);
void launch_resolve(const float *accum, const float *depth_in, int n, int spp, int srgb, float *rgb_mean, uint8_t *rgb8, float *depth_out,
                    uint8_t *sample_count, uint8_t *mask, cudaStream_t st);
size_t trace_rays_scratch_bytes(size_t n);  // device scratch the two ray-buffer launchers need
void launch_trace_closest_rays(const DScene &S, const qgRay *rays, size_t n, qgHit *out, unsigned long long *stats, void *scratch,
                               cudaStream_t st);
void launch_trace_any_rays(const DScene &S, const float *rays6, const float *tmax, size_t n, uint8_t *occ, unsigned long long *stats,
                           void *scratch, cudaStream_t st);
// the any-hit batch with its scratch in a static device array (no allocation; qg_trace_any_noalloc): n <= QB_NOALLOC_MAX_RAYS
#define QB_NOALLOC_MAX_RAYS (1 << 20)
#define QB_NOALLOC_MAX_DEVICES 64
cudaError_t launch_trace_any_rays_noalloc(const DScene &S, const float *rays6, const float *tmax, size_t n, uint8_t *occ,
                                          unsigned long long *stats, cudaStream_t st);
void launch_primary_rays(const WfParams &P, size_t n, qgRay *out, cudaStream_t st);
bool trace_stack_fits(int entries, int extra);  // `extra` more stack entries per lane cost no resident block of the traversal kernels
int wf_trace_grid();   // persistent grid sizes (blocks), from the occupancy calculator
int wf_shadow_grid();
int wf_shade_grid();
int wf_shade_grid_max();  // upper bound of every shade kernel's grid (all material classes): sizes the dead-entry padding of the queues
// SM count of the current device (grid sizes of the non-persistent helpers are multiples of it)
inline int qb_sms()
{
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms > 0 ? sms : 1;
}
void upload_texture_taps(const float *xy64);  // 32 (x,y) tap offsets of Texture::Sample's elliptical filter


}  // namespace qb
