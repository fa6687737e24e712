// shade_kernels.cu — material evaluation, BSDF sampling, next-event estimation and frame resolve (sm_100a).
//
//   k_wf_shade   MtlBlinn_PhotonMap::Shade with photon mapping off (src/materials/MtlBlinn_PhotonMap.cpp:256-500),
//                the material class the reference compiles in (src/materials/materials.h:57), turned from a
//                recursion into one step of a throughput-carrying path:
//                    radiance += T * (emission + sum_lights direct)          (direct via the shadow queue)
//                    T'        = T * BxDF  (x exp(-absorption * z) when the next hit is a back face, :246-248)
//                MultiMtl dispatch (materials.h:70-76), lights (src/lights/lights.cpp:18-144),
//                textures (src/core/texture.cpp:32-114, src/textures/texture.cpp:97-137),
//                sampling distributions (src/core/sampler.cpp:42-120, src/math/math.cpp:37-46).
//   k_resolve    mean, sRGB, clamp, round to 8 bit (src/renderers/renderer.cpp:34-39,348-365).
//
// Random numbers: a counter-based stream per (pixel, sample) drawn in the order the reference's recursion draws
// them (lobe selection, then the lobe's direction sample, then the next hit...), so the same stream fed to the
// reference gives the same path.
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include "wavefront.h"
#include "trace.cuh"  // Ray3
#include "wave_rays.cuh"  // primary_ray, HitRec, finish_hit (the fused kernel computes the hit attributes itself)

namespace qb {

__constant__ float2 c_taps[32];

void upload_texture_taps(const float *xy64) { cudaMemcpyToSymbol(c_taps, xy64, sizeof(float2) * 32); }

struct Rng {
  uint32_t key, n;
  __device__ __forceinline__ float next() { return rng_float(key, n++); }
};

__device__ __forceinline__ V3 ld3(const float *p) { return v3(p[0], p[1], p[2]); }

// ---- textures ---------------------------------------------------------------------------------------------------
// (int) x as the reference's x86-64 build evaluates it (cvttss2si): out-of-range and NaN give INT_MIN, where CUDA's
// conversion would saturate. Differential rays that leave a refracting object almost parallel to a textured plane
// produce texture-space footprints beyond 2^31, so the difference is visible (example_project7_object through glass).
__device__ __forceinline__ int cvt_int(float x) { return fabsf(x) < 2147483648.f ? (int) x : (int) 0x80000000; }

__device__ __forceinline__ V3 tile_clamp(V3 uvw)  // Texture::TileClamp, core/texture.cpp:53-63
{
  V3 u = v3(uvw.x - (float) cvt_int(uvw.x), uvw.y - (float) cvt_int(uvw.y), uvw.z - (float) cvt_int(uvw.z));
  if (u.x < 0) u.x += 1;
  if (u.y < 0) u.y += 1;
  if (u.z < 0) u.z += 1;
  return u;
}

__device__ __forceinline__ V3 texel(const DScene &S, const DTexture &t, int x, int y)
{
  const uchar4 c = __ldg(&S.texels[t.texel_base + (uint32_t) y * (uint32_t) t.width + (uint32_t) x]);
  return v3(c.x / 255.0f, c.y / 255.0f, c.z / 255.0f);
}

__device__ V3 texture_fetch(const DScene &S, const DTexture &t, V3 uvw)
{
  if (t.type == QG_TEX_CHECKER) {  // TextureChecker::Sample, textures/texture.cpp:129-137
    const V3 u = tile_clamp(uvw);
    const V3 c1 = ld3(t.c1), c2 = ld3(t.c2);
    if (u.x <= 0.5f) return u.y <= 0.5f ? c1 : c2;
    return u.y <= 0.5f ? c2 : c1;
  }
  if (t.type != QG_TEX_IMAGE || t.width + t.height == 0) return v3(0.f);
  // TextureFile::Sample, textures/texture.cpp:97-125: v flipped, tiled, bilinear
  const V3 u = tile_clamp(v3(uvw.x, 1.f - uvw.y, uvw.z));
  const int width = t.width, height = t.height;
  const float x = width * u.x;
  const float y = height * u.y;
  int ix = cvt_int(x), iy = cvt_int(y);
  const float fx = x - ix, fy = y - iy;
  if (ix < 0) ix -= (ix / width - 1) * width;
  if (ix >= width) ix -= (ix / width) * width;
  int ixp = ix + 1;
  if (ixp >= width) ixp -= width;
  if (iy < 0) iy -= (iy / height - 1) * height;
  if (iy >= height) iy -= (iy / height) * height;
  int iyp = iy + 1;
  if (iyp >= height) iyp -= height;
  return texel(S, t, ix, iy) * ((1 - fx) * (1 - fy)) + texel(S, t, ixp, iy) * (fx * (1 - fy)) + texel(S, t, ix, iyp) * ((1 - fx) * fy) +
         texel(S, t, ixp, iyp) * (fx * fy);
}

__device__ __forceinline__ V3 texmap_to(const DTexColor &tc, V3 p)  // TextureMap's Transformation::TransformTo
{
  if (tc.identity) return p;
  const V3 a = p - ld3(tc.pos);
  return v3(tc.itm[0] * a.x + tc.itm[3] * a.y + tc.itm[6] * a.z, tc.itm[1] * a.x + tc.itm[4] * a.y + tc.itm[7] * a.z,
            tc.itm[2] * a.x + tc.itm[5] * a.y + tc.itm[8] * a.z);
}

// TexturedColor::Sample(uvw) — single tap (background, environment)
__device__ V3 texcolor_sample1(const DScene &S, const DTexColor &tc, V3 uvw)
{
  const V3 color = ld3(tc.color);
  if (tc.texture < 0) return color;
  return color * texture_fetch(S, S.textures[tc.texture], texmap_to(tc, uvw));
}

// Sample(hInfo, TexturedColor) of MtlBlinn_PhotonMap.cpp:34-39 -> TexturedColor::Sample(uvw, duvw, elliptic=true)
// -> TextureMap::Sample -> Texture::Sample: 1 tap, or 32 taps on the ellipse spanned by duvw when duvw != 0
__device__ V3 texcolor_sample(const DScene &S, const DTexColor &tc, bool has_texture, V3 uvw, V3 d0, V3 d1)
{
  const V3 color = ld3(tc.color);
  if (tc.texture < 0 || !has_texture) return color;
  const DTexture &t = S.textures[tc.texture];
  const V3 u = texmap_to(tc, uvw);
  const V3 e0 = texmap_to(tc, d0 + uvw) - u;
  const V3 e1 = texmap_to(tc, d1 + uvw) - u;
  V3 c = texture_fetch(S, t, u);
  if (length2(e0) + length2(e1) == 0) return color * c;
  for (int i = 1; i < 32; ++i) {
    const float2 xy = c_taps[i];
    c += texture_fetch(S, t, u + xy.x * e0 + xy.y * e1);
  }
  return color * (c / 32.f);
}

__device__ V3 sample_environment(const DScene &S, V3 dir)  // TexturedColor::SampleEnvironment, core/texture.cpp:107-114
{
  const float z = asinf(-dir.z) / 3.14159274101257324f + 0.5f;
  const float x = dir.x / (qabs(dir.x) + qabs(dir.y));
  const float y = dir.y / (qabs(dir.x) + qabs(dir.y));
  return texcolor_sample1(S, S.environment, v3(0.5f, 0.5f, 0.0f) + z * (x * v3(0.5f, 0.5f, 0.f) + y * v3(-0.5f, 0.5f, 0.f)));
}

// ---- sampling distributions --------------------------------------------------------------------------------------
__device__ __forceinline__ V3 to_local_frame(V3 N, V3 sample)  // TransformToLocalFrame, math/math.cpp:37-46
{
  const V3 Z = N;
  const V3 Y = (qabs(Z.x) > qabs(Z.y)) ? normalize(v3(Z.z, 0, -Z.x)) : normalize(v3(0, -Z.z, Z.y));
  const V3 X = normalize(cross(Y, Z));
  const V3 unit = normalize(sample);
  return unit.x * X + unit.y * Y + unit.z * Z;
}
__device__ __forceinline__ V3 cos_weighted_hemisphere(Rng &rng)  // core/sampler.cpp:91-107
{
  const float r1 = rng.next(), r2 = rng.next();
  const float cosTheta = sqrtf(r1);
  const float sinTheta = sqrtf(1 - r1);
  const float phi = 2 * 3.14159274101257324f * r2;
  return v3(sinTheta * cosf(phi), sinTheta * sinf(phi), cosTheta);
}
// Sampler::UniformBall (core/sampler.cpp:42-53): rejection from the cube; z reuses r2 as the reference does.
// The reference loops without bound; 64 rounds leave a miss probability below 1e-20.
__device__ __forceinline__ V3 uniform_ball(Rng &rng, float radius)
{
  V3 p = v3(0.f);
  for (int it = 0; it < 64; ++it) {
    const float r1 = rng.next(), r2 = rng.next(), r3 = rng.next();
    (void) r3;
    p = v3((2.f * r1 - 1.f) * radius, (2.f * r2 - 1.f) * radius, (2.f * r2 - 1.f) * radius);
    if (!(length(p) > radius)) break;
  }
  return p;
}

// SpotLight::GetAttenuation, lights.cpp:128-144
__device__ __forceinline__ float spot_attenuation(const DLight &l, V3 dir)
{
  const float cosTheta = dot(dir, ld3(l.direction));
  if (cosTheta < 0) return 0.f;
  const float r = sqrtf(1.f - cosTheta * cosTheta) / cosTheta;
  if (r > l.outer) return 0.f;
  if (r < l.inner) return 1.f;
  const float q = (l.outer - r) / (l.outer - l.inner);
  return q * q;  // POW(x, 2.f)
}

enum { SEL_TRANSMIT = 0, SEL_REFLECT = 1, SEL_DIFFUSE = 2, SEL_ABSORB = 3 };

__device__ __forceinline__ void accum_add(float *accum, int slot, V3 c)
{
  float *a = accum + 4ll * slot;
  atomicAdd(a + 0, c.x);
  atomicAdd(a + 1, c.y);
  atomicAdd(a + 2, c.z);
}

// stream key of a child path / of an area-light group: (parent key, parent counter, index)
__device__ __forceinline__ uint32_t child_key(uint32_t key, uint32_t n, uint32_t c)
{
  return rng_mix(key ^ rng_mix(n * 0x9E3779B9u + c * 0x85EBCA77u + 0x2545F491u));
}

// warp-converged: lanes with `push` append one (hit, area light) group; its shadow samples are generated by k_wf_area_*
__device__ __forceinline__ void push_area_group(const WfBuffers &B, int d, int lane, uint32_t lt_mask, bool push, V3 p, int light, int slot,
                                                V3 contrib, uint32_t gkey)
{
  const uint32_t mask = __ballot_sync(0xFFFFFFFFu, push);
  if (!mask) return;
  uint32_t qbase = 0;
  if (lane == 0) qbase = atomicAdd(&B.counters[QB_CNT_AREA + d], (uint32_t) __popc(mask));
  qbase = __shfl_sync(0xFFFFFFFFu, qbase, 0);
  if (!push) return;
  const uint32_t q = qbase + __popc(mask & lt_mask);
  if (q >= B.area_capacity) return;
  B.a0[q] = make_float4(p.x, p.y, p.z, __uint_as_float(gkey));
  B.a1[q] = make_float4(__int_as_float(light), __int_as_float(slot), 0.f, 0.f);
  B.a2[q] = make_float4(contrib.x, contrib.y, contrib.z, 0.f);
}

// Output reservation of a shade warp. Every push used to cost one atomic on the queue's counter — one address for the whole
// grid, and the SASS page showed the kernel waiting on exactly those (the SHFL after each leader atomic). A warp now reserves
// QB_OUT_CHUNK slots at a time and hands them out itself; a push that does not fit takes the tail of the old reservation and
// the head of a new one, so the only unused slots are each warp's last tail, which it fills with dead entries at the end
// (path: slot -1, shadow: tmax -1; the consumers skip them, QB_CNT_HOLES_* keeps the ray statistics exact).
struct OutRes {
  uint32_t next, end;
};
__device__ __forceinline__ uint32_t res_take(OutRes &R, uint32_t *counter, int lane, uint32_t m, uint32_t rank)
{
  const uint32_t left = R.end - R.next;
  if (m > left) {
    uint32_t nb = 0;
    if (lane == 0) nb = atomicAdd(counter, (uint32_t) QB_OUT_CHUNK);
    nb = __shfl_sync(0xFFFFFFFFu, nb, 0);
    const uint32_t q = rank < left ? R.next + rank : nb + (rank - left);
    R.next = nb + (m - left);
    R.end = nb + (uint32_t) QB_OUT_CHUNK;
    return q;
  }
  const uint32_t q = R.next + rank;
  R.next += m;
  return q;
}

// the unused tails of a warp's reservations become dead entries (end of a shade kernel)
__device__ __forceinline__ void res_flush(const WfBuffers &B, int d, int lane, const OutRes &rp, const OutRes &rs)
{
  const int nb = (d & 1) ^ 1;
  for (uint32_t q = rp.next + (uint32_t) lane; q < min(rp.end, B.capacity); q += 32u) {
    B.ps0[nb][q] = make_float4(0.f, 0.f, 0.f, 0.f);
    B.ps1[nb][q] = make_float4(1.f, 0.f, __int_as_float(-1), 0.f);
    B.ps2[nb][q] = make_float4(0.f, 0.f, 0.f, 0.f);
    B.ps3[nb][q] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (uint32_t q = rs.next + (uint32_t) lane; q < min(rs.end, B.shadow_capacity); q += 32u) {
    B.s0[q] = make_float4(0.f, 0.f, 0.f, 0.f);
    B.s1[q] = make_float4(1.f, 0.f, -1.f, 0.f);
    B.s2[q] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  if (lane == 0) {
    if (rp.end > rp.next) atomicAdd(&B.counters[QB_CNT_HOLES_PATHS + d + 1], rp.end - rp.next);
    if (rs.end > rs.next) atomicAdd(&B.counters[QB_CNT_HOLES_SHADOW + d], rs.end - rs.next);
  }
}

// FUSED: the hit attributes (k_wf_finish's h0..h3: 64 bytes written and read back per ray, plus a second read of the ray)
// never leave registers — the kernel reads trace's compact record (32 bytes) and runs finish_hit itself. The attribute
// arithmetic is the same code compiled the same way (this translation unit is built with --fmad=false like the traversal's),
// so p, N, uvw and duvw come out bit-equal to the separate pass.
// Residency: 5 blocks / SM (96 registers, 80 B of spills) for the fused kernel, 4 for the plain one — A/B on the tower, fused:
// 4 / 5 / 6 blocks = 13.68 / 13.36 / 13.43 ms of shade time per step
#ifndef QB_SHADE_FUSED_BLOCKS
#define QB_SHADE_FUSED_BLOCKS 5
#endif
template <bool FUSED>
__global__ void __launch_bounds__(QB_BLOCK, FUSED ? QB_SHADE_FUSED_BLOCKS : 4) k_wf_shade(const DScene S, const WfBuffers B, const WfParams P)
{
  const int d = P.depth;
  const uint32_t n = d == 0 ? P.n_paths : min(B.counters[QB_CNT_PATHS + d], B.capacity);
  const int bounce_count = P.max_bounce - d;
  const int lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int pb = d & 1, nb = pb ^ 1;
  const float color_luma_threshold = 0.00001f;
  OutRes rp = {0u, 0u}, rs = {0u, 0u};  // reservations in the path queue of depth d + 1 and in the shadow queue of depth d
  uint32_t w_next = 0, w_end = 0;  // the warp's reserved slice of the queue: one cursor atomic per QB_SHADE_CHUNK paths
  for (;;) {
    if (w_next >= w_end) {
      uint32_t cbase = 0;
      if (lane == 0) cbase = atomicAdd(&B.counters[QB_CNT_FETCH_SHADE + d], (uint32_t) QB_SHADE_CHUNK);
      cbase = __shfl_sync(0xFFFFFFFFu, cbase, 0);
      if (cbase >= n) break;
      w_next = cbase;
      w_end = min(cbase + (uint32_t) QB_SHADE_CHUNK, n);
    }
    const uint32_t base = w_next;
    w_next += 32u;
    // material-sorted order (qgRenderParams.sort_by_material): queue position -> path index through the permutation
    // k_mat_scatter built; lanes past the end get n (no work)
    const uint32_t idx = base + lane < n ? (P.sort_mat ? B.perm[base + lane] : base + lane) : n;

    bool do_lights = false, alive = false;
    V3 p = v3(0.f), N = v3(0.f), V = v3(0.f), T = v3(1.f), Kd = v3(0.f), Ks = v3(0.f);
    float gloss = 0.f;
    int slot = 0;
    // next path
    V3 next_dir = v3(0.f), next_T = v3(0.f), next_absorb = v3(0.f);
    uint32_t next_flags = 0;
    uint32_t gather = 0u;  // photon-map gathers this hit asks for (bit 30 photon map, bit 31 caustics map)
    Rng rng = {0u, 0u};

    if (idx < n) {
      // every record of the path is requested up front: the three dependent round trips (h0 -> path state -> h1..h3) were
      // the kernel's top stalls once the queue atomics were gone (profiles/README.md)
      float4 h0 = make_float4(0.f, 0.f, 0.f, 0.f), h1 = B.h1[idx], h2 = B.h2[idx], h3 = h0;
      if (!FUSED) { h0 = B.h0[idx]; h3 = B.h3[idx]; }
      float4 sa = make_float4(0.f, 0.f, 0.f, 0.f), sb = sa, sc4 = sa, se = sa;
      if (d > 0) { sa = B.ps0[pb][idx]; sb = B.ps1[pb][idx]; sc4 = B.ps2[pb][idx]; se = B.ps3[pb][idx]; }
      HitRec rec;
      int pi = 0, pj = 0, ps = 0;
      if (FUSED) {
        // h1 / h2 still hold trace's record (trace_kernels.cu put_rec); a padding lane of depth 0 has none (node is only
        // looked at for valid paths, the object index is range-checked for the rest)
        rec = unpack_rec(h1, h2);
        const bool hit = rec.k >= 0 && rec.k < S.num_obj;
        h0.x = hit ? rec.z : QB_BIGFLOAT;
        h0.y = __int_as_float(hit ? __ldg(&S.obj_node[rec.k]) : -1);
      }
      const int node = __float_as_int(h0.y);
      V3 ray_dir;
      V3 absorb_prev = v3(0.f);
      uint32_t flags = 0;
      bool valid = true;
      if (d == 0) {
        int i, j, s;
        valid = wf_decode(P, idx, i, j, s, slot);
        pi = i; pj = j; ps = s;
        if (valid) {
          rng.key = rng_key((uint32_t) (j * P.width + i), (uint32_t) s, P.seed);
          rng.n = wf_dof_draws(P);
          atomicAdd(P.accum + 4ll * slot + 3, 1.0f);
          if (P.depth_plane && (s == 0 || P.mode == 1)) P.depth_plane[slot] = node >= 0 ? h0.x : QB_BIGFLOAT;
          if (node < 0) {  // background by screen position (renderer.cpp:336-338)
            const float tx = halton(s, 11) + (float) i, ty = halton(s, 13) + (float) j;
            accum_add(P.accum, slot, texcolor_sample1(S, S.background, v3(tx / (float) P.width, ty / (float) P.height, 0.f)));
          } else {
            const float tx = halton(s, 11) + (float) i, ty = halton(s, 13) + (float) j;
            const V3 cpt = P.scr_a + tx * P.scr_u + ty * P.scr_v;
            ray_dir = normalize(cpt - wf_campos(P, i, j, s));
          }
        }
      } else {
        const float4 a = sa, b = sb, c = sc4, e = se;
        ray_dir = v3(a.w, b.x, b.y);
        slot = __float_as_int(b.z);
        rng.key = __float_as_uint(b.w);
        T = v3(c.x, c.y, c.z);
        const uint32_t packed = __float_as_uint(c.w);
        rng.n = packed >> 8;
        flags = packed & 0xFFu;
        absorb_prev = v3(e.x, e.y, e.z);
        valid = slot >= 0;  // slot -1: a dead entry (unused tail of a shade warp's reservation)
        if (node < 0 && valid) accum_add(P.accum, slot, T * sample_environment(S, ray_dir));  // ComputeSecondaryRay miss (:250)
      }
      if (valid && node >= 0) {
        if (FUSED) {  // k_wf_finish's body (trace_kernels.cu): the ray again, then the attributes of the recorded hit
          Ray3 c, x, y;
          if (d == 0) {
            primary_ray(P, pi, pj, ps, c, x, y);
          } else {
            c.p = v3(sa.x, sa.y, sa.z);
            c.d = v3(sa.w, sb.x, sb.y);
            x = c;
            y = c;
          }
          HitAttr h;
          finish_hit(S, rec, c, x, y, d == 0, P.stale_tex != 0, h, nullptr);
          h0.z = __int_as_float(h.mtl_id);
          h0.w = __int_as_float((h.front ? 1 : 0) | (h.has_texture ? 2 : 0));
          h1 = make_float4(h.p.x, h.p.y, h.p.z, h.N.x);
          h2 = make_float4(h.N.y, h.N.z, h.uvw.x, h.uvw.y);
          h3 = make_float4(h.duvw0.x, h.duvw0.y, h.duvw1.x, h.duvw1.y);
        }
        const float z = h0.x;
        const int mtl_id = __float_as_int(h0.z);
        const int hflags = __float_as_int(h0.w);
        const bool front = hflags & 1, has_tex = hflags & 2;
        p = v3(h1.x, h1.y, h1.z);
        N = v3(h1.w, h2.x, h2.y);
        const V3 uvw = v3(h2.z, h2.w, 0.f);
        const V3 d0 = v3(h3.x, h3.y, 0.f), d1 = v3(h3.z, h3.w, 0.f);
        // Beer-Lambert over the segment that ended on a back face (MtlBlinn_PhotonMap.cpp:246-248, materials.cpp:32-38)
        if (d > 0 && !front) T = T * v3(expf(-absorb_prev.x * z), expf(-absorb_prev.y * z), expf(-absorb_prev.z * z));
        int mi = S.nodes[node].material;
        bool white = false;
        if (mi >= 0 && S.materials[mi].kind == QG_MTL_MULTI) {  // MultiMtl::Shade
          const DMaterial &mm = S.materials[mi];
          if (mtl_id >= 0 && mtl_id < mm.sub_count) mi = S.multi[mm.sub_first + mtl_id];
          else white = true;
        }
        if (white) {
          accum_add(P.accum, slot, T);
        } else if (mi >= 0) {
          const DMaterial &M = S.materials[mi];
          V = -ray_dir;
          V3 color = texcolor_sample(S, M.emission, has_tex, uvw, d0, d1);
          // ComputeFresnel (:65-105)
          const V3 Y = dot(N, V) > 0.f ? N : -N;
          const V3 Zf = cross(V, Y);
          const V3 X = normalize(cross(Y, Zf));
          const float nIOR = front ? 1.f / M.ior : M.ior;
          const float cosI = dot(N, V);
          const float sinI = sqrtf(1 - cosI * cosI);
          const float sinO = qmax(0.f, qmin(1.f, sinI * nIOR));
          const float cosO = sqrtf(1.f - sinO * sinO);
          const V3 tDir = -X * sinO - Y * cosO;
          const V3 rDir = 2.f * N * (dot(N, V)) - V;
          const bool total_reflection = (nIOR * sinI) > 1.001f;
          const float C = (nIOR - 1.f) * (nIOR - 1.f) / ((nIOR + 1.f) * (nIOR + 1.f));
          const float rC = C + (1.f - C) * powf(1.f - qabs(cosI), 5.f);
          const float tC = 1.f - rC;
          const V3 tK = texcolor_sample(S, M.refraction, has_tex, uvw, d0, d1);
          const V3 rK = texcolor_sample(S, M.reflection, has_tex, uvw, d0, d1);
          const V3 sT = total_reflection ? v3(0.f) : tK * tC;
          const V3 sR = total_reflection ? (rK + tK) : (rK + tK * rC);
          Ks = texcolor_sample(S, M.specular, has_tex, uvw, d0, d1);
          Kd = texcolor_sample(S, M.diffuse, has_tex, uvw, d0, d1);
          gloss = M.gloss;
          // RandomSelectMtl (:107-150)
          const float lumaT = luma(sT), lumaR = luma(sR), lumaD = luma(Kd);
          const float r = rng.next();
          const float coefT = lumaT, coefR = coefT + lumaR, coefD = coefR + lumaD, coefSum = coefD + M.kill;
          const float select = r * coefSum;
          int sel;
          if (select < coefT && lumaT > color_luma_threshold) sel = SEL_TRANSMIT;
          else if (select < coefR && lumaR > color_luma_threshold) sel = SEL_REFLECT;
          else if (select < coefD && lumaD > color_luma_threshold) sel = SEL_DIFFUSE;
          else sel = SEL_ABSORB;
          // Scene::usePhotonMap (:341-352): a diffuse pick gathers the caustics map, and the photon map instead of a further
          // diffuse bounce once the path has had one; the bounce rule below is the same with and without the maps
          if (P.use_photon_map && sel == SEL_DIFFUSE) gather = 0x80000000u | ((flags & QB_FLAG_DIFFUSE_HIT) ? 0x40000000u : 0u);
          if (bounce_count > 0) {
            if (sel == SEL_REFLECT) {  // SampleReflectionBxDF (:179-203)
              V3 dir = rDir;
              if (M.refl_gloss > 0.f) {
                for (int it = 0; it < 64; ++it) {
                  dir = normalize(normalize(rDir) + uniform_ball(rng, 2.f * M.refl_gloss));
                  if (!(dot(dir, Y) < 0)) break;
                }
              }
              alive = true; next_dir = dir; next_T = T * sR;
            } else if (sel == SEL_TRANSMIT) {  // SampleTransmitBxDF (:152-178)
              V3 dir = tDir;
              if (M.refr_gloss > 0.f) {
                for (int it = 0; it < 64; ++it) {
                  dir = normalize(normalize(tDir) + uniform_ball(rng, 2.f * M.refr_gloss));
                  if (!(dot(dir, Y) > 0)) break;
                }
              }
              alive = true; next_dir = dir; next_T = T * sT;
            } else if (sel == SEL_DIFFUSE && !(flags & QB_FLAG_DIFFUSE_HIT) && front) {  // SampleDiffuseBxDF (:204-224)
              const V3 dir = to_local_frame(N, cos_weighted_hemisphere(rng));
              const V3 L = normalize(dir);
              const V3 H = normalize(V + L);
              const float cosNH = qmax(0.f, dot(N, H));
              alive = true; next_dir = dir; next_T = T * (Kd + Ks * powf(cosNH, gloss));
              next_flags = QB_FLAG_DIFFUSE_HIT;
            }
          }
          next_absorb = ld3(M.absorption);
          accum_add(P.accum, slot, T * color);
          do_lights = true;
        }
      }
    }

    // ---- next-event estimation: one shadow ray per non-ambient light (:484-498), warp-aggregated queue push
    const float norm_di = S.num_lights > 0 ? 1.f / (float) S.num_lights : 1.f;
    for (int l = 0; l < S.num_lights; ++l) {
      const DLight &lt = S.lights[l];
      if (lt.type == QG_LIGHT_AMBIENT) continue;  // uniform across the warp
      const bool area = lt.type != QG_LIGHT_DIRECT && lt.size > 0.01f;  // lights.cpp:52,87 (uniform across the warp)
      bool push = false;
      V3 sdir = v3(0.f), contrib = v3(0.f);
      float tmax = QB_BIGFLOAT;
      if (do_lights) {
        V3 I = ld3(lt.intensity), Ldir;
        if (lt.type == QG_LIGHT_DIRECT) {
          sdir = normalize(-ld3(lt.direction));
          Ldir = ld3(lt.direction);
        } else {
          const V3 dv = ld3(lt.position) - p;
          sdir = normalize(dv);
          tmax = length(dv);
          if (!area) I = I * qmin(1.f, 1.f / length2(dv));  // InverseSquareFalloff (an area light folds it into its shadow estimate)
          Ldir = normalize(p - ld3(lt.position));
          if (lt.type == QG_LIGHT_SPOT) I = I * spot_attenuation(lt, Ldir);
        }
        const V3 L = normalize(-Ldir);
        const V3 H = normalize(V + L);
        const float cosNL = qmax(0.f, dot(N, L)), cosNH = qmax(0.f, dot(N, H));
        contrib = T * ((I * norm_di) * cosNL * (Kd + Ks * powf(cosNH, gloss)));
        push = (contrib.x != 0.f) || (contrib.y != 0.f) || (contrib.z != 0.f);
      }
      if (area) {
        push_area_group(B, d, lane, lt_mask, push, p, l, slot, contrib, child_key(rng.key, rng.n, 0x100u + (uint32_t) l));
        continue;
      }
      const uint32_t mask = __ballot_sync(0xFFFFFFFFu, push);
      if (mask) {
        const uint32_t q = res_take(rs, &B.counters[QB_CNT_SHADOW + d], lane, (uint32_t) __popc(mask), (uint32_t) __popc(mask & lt_mask));
        if (push && q < B.shadow_capacity) {
          B.s0[q] = make_float4(p.x, p.y, p.z, sdir.x);
          B.s1[q] = make_float4(sdir.y, sdir.z, tmax, __int_as_float(slot));
          B.s2[q] = make_float4(contrib.x, contrib.y, contrib.z, 0.f);
        }
      }
    }

    // ---- photon-map gathers (:426-458) are queued for k_photon_gather: the k-nearest search keeps 100 candidates per thread
    if (P.use_photon_map) {
      const uint32_t gm = __ballot_sync(0xFFFFFFFFu, gather != 0u);
      if (gm) {
        uint32_t gb = 0;
        if (lane == 0) gb = atomicAdd(&B.counters[QB_CNT_GATHER + d], (uint32_t) __popc(gm));
        gb = __shfl_sync(0xFFFFFFFFu, gb, 0);
        const uint32_t q = gb + (uint32_t) __popc(gm & lt_mask);
        if (gather && q < B.capacity) {
          B.g0[q] = make_float4(p.x, p.y, p.z, N.x);
          B.g1[q] = make_float4(N.y, N.z, V.x, V.y);
          B.g2[q] = make_float4(V.z, Kd.x, Kd.y, Kd.z);
          B.g3[q] = make_float4(Ks.x, Ks.y, Ks.z, gloss);
          B.g4[q] = make_float4(T.x, T.y, T.z, __uint_as_float((uint32_t) slot | gather));
        }
      }
    }

    // ---- queue compaction: surviving paths go to the state buffer of depth d+1
    const uint32_t amask = __ballot_sync(0xFFFFFFFFu, alive);
    if (amask) {
      const uint32_t q = res_take(rp, &B.counters[QB_CNT_PATHS + d + 1], lane, (uint32_t) __popc(amask), (uint32_t) __popc(amask & lt_mask));
      if (alive && q < B.capacity) {
        const V3 nd = normalize(next_dir);  // DiffRay::Normalize in ComputeSecondaryRay (:235-236)
        B.ps0[nb][q] = make_float4(p.x, p.y, p.z, nd.x);
        B.ps1[nb][q] = make_float4(nd.y, nd.z, __int_as_float(slot), __uint_as_float(rng.key));
        B.ps2[nb][q] = make_float4(next_T.x, next_T.y, next_T.z, __uint_as_float((rng.n << 8) | next_flags));
        B.ps3[nb][q] = make_float4(next_absorb.x, next_absorb.y, next_absorb.z, 0.f);
      }
    }
  }
  res_flush(B, d, lane, rp, rs);
}

}  // namespace qb
#include "shade_alt.cuh"
#include "photon.cuh"
namespace qb {

// ---- resolve ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float linear_to_srgb(float c)  // renderer.cpp:34-39
{
  const float a = 0.055f;
  if (c < 0.0031308f) return 12.92f * c;
  return (1.f + a) * powf(c, 1.f / 2.4f) - a;
}

__global__ void k_resolve(const float *__restrict__ accum, const float *__restrict__ depth_in, int n, int spp, int srgb,
                          float *__restrict__ rgb_mean, uint8_t *__restrict__ rgb8, float *__restrict__ depth_out,
                          uint8_t *__restrict__ sample_count, uint8_t *__restrict__ mask)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 a = reinterpret_cast<const float4 *>(accum)[i];
  const float w = a.w;
  V3 c = w > 0.f ? v3(a.x / w, a.y / w, a.z / w) : v3(0.f);
  if (rgb_mean) { rgb_mean[3 * i] = c.x; rgb_mean[3 * i + 1] = c.y; rgb_mean[3 * i + 2] = c.z; }
  if (rgb8) {
    if (srgb) c = v3(linear_to_srgb(c.x), linear_to_srgb(c.y), linear_to_srgb(c.z));
    c = v3(qmax(0.f, qmin(1.f, c.x)), qmax(0.f, qmin(1.f, c.y)), qmax(0.f, qmin(1.f, c.z)));
    rgb8[3 * i + 0] = (uint8_t) roundf(c.x * 255.f);
    rgb8[3 * i + 1] = (uint8_t) roundf(c.y * 255.f);
    rgb8[3 * i + 2] = (uint8_t) roundf(c.z * 255.f);
  }
  if (depth_out) depth_out[i] = depth_in ? depth_in[i] : 0.f;
  // sampleCountBuffer = 255 * samples taken / sppMax (renderer.cpp:362-363); mask = pixel rendered (:364)
  if (sample_count) sample_count[i] = (uint8_t) (255.f * w / (float) spp);
  if (mask) mask[i] = w > 0.f ? 1 : 0;
}

// ---- material-sorted shading (optional, qgRenderParams.sort_by_material) -----------------------------------------------
// Counting sort of the depth's shade queue by the material that will shade each hit (MultiMtl resolved), so the lanes of
// a warp evaluate the same textures / lobes. Three small kernels: histogram, scan, scatter -> B.perm.
__device__ __forceinline__ uint32_t shade_key(const DScene &S, const WfBuffers &B, uint32_t idx)
{
  const float4 h0 = B.h0[idx];
  const int node = __float_as_int(h0.y);
  if (node < 0) return 0u;
  int mi = S.nodes[node].material;
  if (mi >= 0 && S.materials[mi].kind == QG_MTL_MULTI) {
    const DMaterial &mm = S.materials[mi];
    const int mtl_id = __float_as_int(h0.z);
    mi = (mtl_id >= 0 && mtl_id < mm.sub_count) ? S.multi[mm.sub_first + mtl_id] : -1;
  }
  return (uint32_t) min(mi + 2, QB_MAT_BINS - 1);
}
__global__ void k_mat_hist(const DScene S, const WfBuffers B, const WfParams P)
{
  __shared__ uint32_t s_bins[QB_MAT_BINS];
  for (int i = threadIdx.x; i < QB_MAT_BINS; i += blockDim.x) s_bins[i] = 0u;
  __syncthreads();
  const uint32_t n = P.depth == 0 ? P.n_paths : min(B.counters[QB_CNT_PATHS + P.depth], B.capacity);
  for (uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) atomicAdd(&s_bins[shade_key(S, B, idx)], 1u);
  __syncthreads();
  for (int i = threadIdx.x; i < QB_MAT_BINS; i += blockDim.x)
    if (s_bins[i]) atomicAdd(&B.mat_bins[i], s_bins[i]);
}
__global__ void k_mat_scan(const WfBuffers B)
{
  if (threadIdx.x != 0) return;
  uint32_t run = 0u;
  for (int i = 0; i < QB_MAT_BINS; ++i) {
    B.mat_bins[QB_MAT_BINS + i] = run;
    run += B.mat_bins[i];
    B.mat_bins[i] = 0u;  // clean for the next depth
  }
}
// Tile-local counting scatter: a block ranks a tile of QB_SORT_TILE hits per material in shared memory, reserves one range
// per non-empty bin with ONE global atomic, and writes its indices there — the global atomics drop from one per hit to
// one per (tile, material), and a material's run in B.perm is made of chunks whose indices come from one 1024-hit window
// of the queue, so the gathered path-state reads of shade stay sector-local.
#define QB_SORT_ITEMS 4
#define QB_SORT_TILE (256 * QB_SORT_ITEMS)
__global__ void __launch_bounds__(256) k_mat_scatter(const DScene S, const WfBuffers B, const WfParams P)
{
  __shared__ uint32_t s_cnt[QB_MAT_BINS], s_base[QB_MAT_BINS];
  const uint32_t n = P.depth == 0 ? P.n_paths : min(B.counters[QB_CNT_PATHS + P.depth], B.capacity);
  for (uint32_t tile = blockIdx.x * QB_SORT_TILE; tile < n; tile += gridDim.x * QB_SORT_TILE) {
    s_cnt[threadIdx.x] = 0u;  // blockDim.x == QB_MAT_BINS == 256
    __syncthreads();
    uint32_t key[QB_SORT_ITEMS], rank[QB_SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < QB_SORT_ITEMS; ++i) {
      const uint32_t idx = tile + i * 256 + threadIdx.x;
      key[i] = idx < n ? shade_key(S, B, idx) : 0u;
      rank[i] = idx < n ? atomicAdd(&s_cnt[key[i]], 1u) : 0u;
    }
    __syncthreads();
    if (s_cnt[threadIdx.x]) s_base[threadIdx.x] = atomicAdd(&B.mat_bins[QB_MAT_BINS + threadIdx.x], s_cnt[threadIdx.x]);
    __syncthreads();
#pragma unroll
    for (int i = 0; i < QB_SORT_ITEMS; ++i) {
      const uint32_t idx = tile + i * 256 + threadIdx.x;
      if (idx < n) B.perm[s_base[key[i]] + rank[i]] = idx;
    }
    __syncthreads();
  }
}
void launch_mat_sort(const DScene &S, const WfBuffers &B, const WfParams &P, cudaStream_t st)
{
  k_mat_hist<<<qb_sms() * 4, 256, 0, st>>>(S, B, P);
  k_mat_scan<<<1, 32, 0, st>>>(B);
  static_assert(QB_MAT_BINS == 256, "k_mat_scatter uses one thread per bin");
  k_mat_scatter<<<qb_sms() * 8, 256, 0, st>>>(S, B, P);
}

// ---- adaptive sampling: SuperSamplerHalton (src/scene/scene.cpp:83-123) -----------------------------------------------------
// The reference loops per pixel: sample, Accumulate (running mean + a variance proxy), stop when s >= sppMin and every
// channel's proxy is below (0.005, 0.001, 0.005), or at sppMax. Sample s of a pixel is a pure function of (pixel, s, seed)
// here, so pixels still running get their next `chunk` samples traced together (one path each, radiance kept per path),
// and k_adaptive_update replays the reference's loop over them in order; samples past the stop are discarded unseen.
__device__ __forceinline__ bool adaptive_loop(int s, V3 sd, int spp_min, int spp_max)  // SuperSamplerHalton::Loop
{
  return s < spp_min || (s >= spp_min && s < spp_max && (sd.x > 0.005f || sd.y > 0.001f || sd.z > 0.005f));
}

// pixels of this tile shard that are still running -> compact list + their next `chunk` (i, j, s) triples
__global__ void k_adaptive_build(int W, int H, int tile_rank, int tile_ranks, const int32_t *__restrict__ count,
                                 const uint8_t *__restrict__ active, int chunk, uint32_t *__restrict__ n_active, int32_t *__restrict__ list,
                                 int32_t *__restrict__ triples)
{
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= W * H || !active[pix]) return;
  const int i = pix % W, j = pix / W;
  const int tiles_x = (W + 3) / 4;
  const int k = (j / 4) * tiles_x + (i / 4);  // 4x4 tile index; the shard owns k = rank (mod ranks) (renderer.cpp:383-398)
  if (k % tile_ranks != tile_rank) return;
  const uint32_t a = atomicAdd(n_active, 1u);
  list[a] = pix;
  const int s0 = count[pix];
  for (int c = 0; c < chunk; ++c) {
    int32_t *t = triples + 3ll * ((long long) a * chunk + c);
    t[0] = i; t[1] = j; t[2] = s0 + c;
  }
}

// SuperSamplerHalton::Accumulate over the chunk, in sample order; stops where Loop() stops
__global__ void k_adaptive_update(uint32_t n_active, const int32_t *__restrict__ list, int chunk, const float4 *__restrict__ rad,
                                  const float *__restrict__ rad_depth, int spp_min, int spp_max, int32_t *__restrict__ count,
                                  uint8_t *__restrict__ active, float4 *__restrict__ mean, float4 *__restrict__ sdev, float *__restrict__ depth)
{
  const uint32_t a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= n_active) return;
  const int pix = list[a];
  int s = count[pix];
  const float4 m4 = mean[pix], s4 = sdev[pix];
  V3 color = v3(m4.x, m4.y, m4.z), sd = v3(s4.x, s4.y, s4.z);
  for (int c = 0; c < chunk && adaptive_loop(s, sd, spp_min, spp_max); ++c) {
    const float4 r = rad[(long long) a * chunk + c];
    if (s == 0 && depth) depth[pix] = rad_depth[(long long) a * chunk + c];
    const V3 dc = (v3(r.x, r.y, r.z) - color) / (float) (s + 1);
    color += dc;
    if (s > 0) sd += dc * dc * (float) (s + 1) - sd / (float) s;
    ++s;
  }
  count[pix] = s;
  mean[pix] = make_float4(color.x, color.y, color.z, 0.f);
  sdev[pix] = make_float4(sd.x, sd.y, sd.z, 0.f);
  active[pix] = adaptive_loop(s, sd, spp_min, spp_max) ? 1 : 0;
}

// mean and sample count -> the {sum, weight} accumulation buffer the resolve step reads
__global__ void k_adaptive_finish(int n, const int32_t *__restrict__ count, const float4 *__restrict__ mean, float *__restrict__ accum)
{
  const int pix = blockIdx.x * blockDim.x + threadIdx.x;
  if (pix >= n) return;
  const float w = (float) count[pix];
  if (w <= 0.f) return;
  const float4 m = mean[pix];
  float4 *a = reinterpret_cast<float4 *>(accum) + pix;
  *a = make_float4(a->x + m.x * w, a->y + m.y * w, a->z + m.z * w, a->w + w);
}

void launch_adaptive_build(int W, int H, int tile_rank, int tile_ranks, const int32_t *count, const uint8_t *active, int chunk,
                           uint32_t *n_active, int32_t *list, int32_t *triples, cudaStream_t st)
{
  k_adaptive_build<<<(W * H + 255) / 256, 256, 0, st>>>(W, H, tile_rank, tile_ranks, count, active, chunk, n_active, list, triples);
}
void launch_adaptive_update(uint32_t n_active, const int32_t *list, int chunk, const float *rad, const float *rad_depth, int spp_min,
                            int spp_max, int32_t *count, uint8_t *active, float *mean, float *sdev, float *depth, cudaStream_t st)
{
  if (n_active == 0) return;
  k_adaptive_update<<<(n_active + 255) / 256, 256, 0, st>>>(n_active, list, chunk, (const float4 *) rad, rad_depth, spp_min, spp_max, count,
                                                            active, (float4 *) mean, (float4 *) sdev, depth);
}
void launch_adaptive_finish(int n, const int32_t *count, const float *mean, float *accum, cudaStream_t st)
{
  k_adaptive_finish<<<(n + 255) / 256, 256, 0, st>>>(n, count, (const float4 *) mean, accum);
}

// wave epilogue: fold the wave's queue counters into the running ray statistics
__global__ void k_wf_stats(const uint32_t *__restrict__ counters, unsigned long long *__restrict__ stats, int max_bounce)
{
  unsigned long long closest = 0, shadow = 0;
  for (int d = 0; d <= max_bounce; ++d) {
    if (d >= 1) closest += counters[QB_CNT_PATHS + d] - counters[QB_CNT_HOLES_PATHS + d];
    shadow += counters[QB_CNT_SHADOW + d] - counters[QB_CNT_HOLES_SHADOW + d] + (unsigned long long) counters[QB_CNT_AREA + d] * QB_AREA_SPP_MAX;
  }
  stats[0] += closest;
  stats[1] += shadow;
}
void launch_wf_stats(const WfBuffers &B, int max_bounce, cudaStream_t st) { k_wf_stats<<<1, 1, 0, st>>>(B.counters, B.stats, max_bounce); }

// persistent grid of one shade kernel: every SM filled to that kernel's own residency (the classes differ a lot in
// registers, so one grid for all would hold the default class at the heaviest class's 3 blocks / SM)
#define QB_SHADE_MAX_BLOCKS 4
int wf_shade_grid_max() { return qb_sms() * std::max(QB_SHADE_MAX_BLOCKS, QB_SHADE_FUSED_BLOCKS); }
static int persistent_grid_shade(const void *kernel, int max_blocks = QB_SHADE_MAX_BLOCKS)
{
  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, QB_BLOCK, 0);
  if (per_sm < 1) per_sm = 1;
  if (per_sm > max_blocks) per_sm = max_blocks;  // measured: 4 blocks / SM is the knee of the plain kernels (xmas shade 104.7 / 92.9 / 93.5 ms at 3 / 4 / 5, tower 11.3 / 10.8 / 13.1)
  if (const char *e = getenv("QB_SHADE_BLOCKS")) per_sm = std::max(1, std::min(per_sm, atoi(e)));  // A/B override
  return sms * per_sm;
}
int wf_shade_grid() { static int g = persistent_grid_shade((const void *) k_wf_shade<false>); return g; }
static int wf_shade_grid_fused() { static int g = persistent_grid_shade((const void *) k_wf_shade<true>, QB_SHADE_FUSED_BLOCKS); return g; }
template <int C, bool PRIMARY> static int alt_grid()
{
  static int g = persistent_grid_shade((const void *) k_wf_shade_alt<C, PRIMARY>);
  return g;
}

void launch_wf_shade(const DScene &S, const WfBuffers &B, const WfParams &P, int grid, cudaStream_t st)
{
#define ALT(C) (P.depth == 0 ? k_wf_shade_alt<C, true><<<alt_grid<C, true>(), QB_BLOCK, 0, st>>>(S, B, P) : k_wf_shade_alt<C, false><<<alt_grid<C, false>(), QB_BLOCK, 0, st>>>(S, B, P))
  switch (P.mtl_class) {
    case QG_CLASS_PATHTRACING: ALT(QG_CLASS_PATHTRACING); break;
    case QG_CLASS_MONTECARLOGI: ALT(QG_CLASS_MONTECARLOGI); break;
    case QG_CLASS_BLINN_BASIC: ALT(QG_CLASS_BLINN_BASIC); break;
    case QG_CLASS_PHONG_BASIC: ALT(QG_CLASS_PHONG_BASIC); break;
    default:
      if (P.fused_finish) k_wf_shade<true><<<wf_shade_grid_fused(), QB_BLOCK, 0, st>>>(S, B, P);
      else k_wf_shade<false><<<grid, QB_BLOCK, 0, st>>>(S, B, P);
      break;
  }
}
void launch_resolve(const float *accum, const float *depth_in, int n, int spp, int srgb, float *rgb_mean, uint8_t *rgb8, float *depth_out,
                    uint8_t *sample_count, uint8_t *mask, cudaStream_t st)
{
  k_resolve<<<(n + 255) / 256, 256, 0, st>>>(accum, depth_in, n, spp, srgb, rgb_mean, rgb8, depth_out, sample_count, mask);
}

}  // namespace qb
