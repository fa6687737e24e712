// photon.cuh — photon / caustics maps on the GPU (included at the end of shade_kernels.cu: it shares the material code).
//
//   emission pass   Renderer::ComputeScene, src/renderers/renderer.cpp:118-291: photons leave the point lights, are traced
//                   with Scene::TraceNodeNormal, stored at diffuse surfaces after their first bounce and continued by
//                   MtlBlinn_PhotonMap::RandomPhotonBounce (src/materials/MtlBlinn_PhotonMap.cpp:503-575)
//   estimate        cyPhotonMap::EstimateIrradiance<100> with the quadratic filter (src/ext/cyPhotonMap.h:333-395, LocatePhotons
//                   :424-495): the 100 nearest photons inside the map's radius whose direction faces the surface
//   gather          MtlBlinn_PhotonMap::Shade, :426-458: I * cos(N, L) * (diffuse + specular * cos(N, H)^gloss)
//
// The reference emits photons one after another from one random stream until the map is full. Here emission e has its own
// counter-based stream (key = rng_key(e, map, seed)), emissions are traced in batches through the ray-buffer kernels of
// trace_kernels.cu, and the host keeps the photons of emissions 0, 1, 2, ... in that order until the map is full — the same
// stopping rule, independent of batch size. The estimate walks cyPhotonMap's own balanced kd-tree (built on the host,
// api_photon.inc) in the reference's order, because its result is order-dependent (photon_estimate below).
#pragma once

namespace qb {

#define QB_PM_K 100  // EstimateIrradiance<100> (MtlBlinn_PhotonMap.cpp:430,445)

__device__ __forceinline__ V3 uniform_sphere(Rng &rng)  // Sampler::UniformSphere, core/sampler.cpp:55-70
{
  float r1 = rng.next();
  const float r2 = rng.next();
  r1 = r1 * 2.f - 1.f;
  const float cosTheta = r1;
  const float sinTheta = sqrtf(1 - r1 * r1);
  const float phi = 2 * 3.14159274101257324f * r2;
  return v3(sinTheta * cosf(phi), sinTheta * sinf(phi), cosTheta);
}
__device__ __forceinline__ V3 uniform_hemisphere(Rng &rng)  // Sampler::UniformHemisphere, core/sampler.cpp:72-86
{
  const float r1 = rng.next(), r2 = rng.next();
  const float cosTheta = r1;
  const float sinTheta = sqrtf(1 - r1 * r1);
  const float phi = 2 * 3.14159274101257324f * r2;
  return v3(sinTheta * cosf(phi), sinTheta * sinf(phi), cosTheta);
}

// ---- irradiance estimate ---------------------------------------------------------------------------------------------------
// LocatePhotons' bookkeeping, restated: candidates are photons with dist2 < dist2[0] whose direction faces the surface
// (dot(dir, N) < 0). The first 100 are kept as they come; the 100th turns the list into a max-heap; from the 101st on the
// farthest is replaced and dist2[0] — the squared search radius, which is also the filter's and the area's radius — shrinks
// to the heap's maximum. So with at most 100 candidates the estimate is normalised by the map's own radius.
struct PhotonHeap {
  float dist2[QB_PM_K + 1];
  uint32_t id[QB_PM_K + 1];
  int found;
};

__device__ __forceinline__ void heap_offer(PhotonHeap &h, float dist2, uint32_t id)
{
  if (h.found < QB_PM_K) {
    h.found++;
    h.dist2[h.found] = dist2;
    h.id[h.found] = id;
    if (h.found == QB_PM_K) {  // build the heap
      const int half_found = h.found >> 1;
      for (int k = half_found; k >= 1; k--) {
        int parent = k;
        const uint32_t tp = h.id[k];
        const float td2 = h.dist2[k];
        while (parent <= half_found) {
          int j = parent + parent;
          if (j < h.found && h.dist2[j] < h.dist2[j + 1]) j++;
          if (td2 >= h.dist2[j]) break;
          h.dist2[parent] = h.dist2[j];
          h.id[parent] = h.id[j];
          parent = j;
        }
        h.id[parent] = tp;
        h.dist2[parent] = td2;
      }
    }
  } else {
    int parent = 1, j = 2;
    while (j <= h.found) {
      if (j < h.found && h.dist2[j] < h.dist2[j + 1]) j++;
      if (dist2 > h.dist2[j]) break;
      h.dist2[parent] = h.dist2[j];
      h.id[parent] = h.id[j];
      parent = j;
      j <<= 1;
    }
    h.id[parent] = id;
    h.dist2[parent] = dist2;
    h.dist2[0] = h.dist2[1];
  }
}

// One photon of the tree against the query (the tail of LocatePhotons, cyPhotonMap.h:443-494)
__device__ __forceinline__ void photon_visit(const DPhotonMap &M, uint32_t idx, V3 pos, V3 N, PhotonHeap &h)
{
  const float4 a4 = M.pa[idx];
  const V3 dif = v3(a4.x, a4.y, a4.z) - pos;
  const float dist2 = length2(dif);
  if (dist2 < h.dist2[0]) {
    const float4 b4 = M.pb[idx];
    if (dot(v3(b4.x, b4.y, b4.z), N) >= 0) return;  // the photon arrived from behind the surface
    heap_offer(h, dist2, idx);
  }
}

// EstimateIrradiance<100>(I, D, radius, pos, &N, 1.f, FILTER_TYPE_QUADRATIC). Returns false when no photon was found
// (I = 0, D = 0 then, as the reference leaves them). LocatePhotons (:424-442) is a recursion — near child, far child if the
// splitting plane is closer than the current search radius, then the node's own photon — replayed here with an explicit
// stack in exactly that order: the 101st candidate replaces the heap's farthest photon even when it is farther itself
// (heap_offer's else branch only ever compared it with the map's radius), so which photons survive depends on the order.
__device__ __noinline__ bool photon_estimate(const DPhotonMap &M, V3 pos, V3 N, V3 &irrad, V3 &direction)
{
  irrad = v3(0.f);
  direction = v3(0.f);
  if (M.count == 0u) return false;
  PhotonHeap h;
  h.found = 0;
  h.dist2[0] = M.radius * M.radius;
  const float q[3] = {pos.x, pos.y, pos.z};
  uint32_t stack[64];  // (index << 2) | stage; the tree is balanced: depth <= 32
  int sp = 0;
  stack[sp++] = 1u << 2;
  while (sp) {
    const uint32_t e = stack[--sp];
    const uint32_t idx = e >> 2, stage = e & 3u;
    if (stage == 2u || (int) idx >= M.half) {
      photon_visit(M, idx, pos, N, h);
      continue;
    }
    const float4 a4 = M.pa[idx];
    const int axis = (int) ((__float_as_uint(M.pb[idx].w) >> 24) & 3u);
    const float p_axis = axis == 0 ? a4.x : (axis == 1 ? a4.y : a4.z);
    const float dist = q[axis] - p_axis;
    if (stage == 0u) {
      stack[sp++] = (idx << 2) | 1u;
      stack[sp++] = (dist > 0 ? 2u * idx + 1u : 2u * idx) << 2;
    } else {
      stack[sp++] = (idx << 2) | 2u;
      if (dist * dist < h.dist2[0]) stack[sp++] = (dist > 0 ? 2u * idx : 2u * idx + 1u) << 2;
    }
  }
  for (int i = 1; i <= h.found; i++) {
    const float4 a4 = M.pa[h.id[i]], b4 = M.pb[h.id[i]];
    const uint32_t c = __float_as_uint(b4.w);  // low 24 bits: the Color24
    const V3 power = v3((float) (c & 255u) / 255.0f, (float) ((c >> 8) & 255u) / 255.0f, (float) ((c >> 16) & 255u) / 255.0f) * a4.w;
    const float filter = 1 - h.dist2[i] / h.dist2[0];
    irrad = irrad + filter * power;
    direction = direction + v3(b4.x, b4.y, b4.z) * (filter * a4.w);
  }
  if (h.found > 0) {
    const float area = (3.14159274101257324f * 0.5f) * h.dist2[0];
    if (area > 0) irrad = irrad * (1.0f / area);
    direction = normalize(direction);
  }
  return h.found > 0;
}

// the gather term of Shade (:426-458) for one map
__device__ __forceinline__ V3 photon_gather_term(const DPhotonMap &M, V3 p, V3 N, V3 V, V3 Kd, V3 Ks, float gloss)
{
  V3 I, D;
  photon_estimate(M, p, N, I, D);
  if (!(luma(I) > 0.00001f)) return v3(0.f);  // "in case we found nothing"
  const V3 L = -normalize(D);
  const V3 H = normalize(V + L);
  const float cosNL = qmax(0.f, dot(N, L));
  const float cosNH = qmax(0.f, dot(N, H));
  return I * cosNL * (Kd + Ks * powf(cosNH, gloss));
}

__global__ void __launch_bounds__(128) k_photon_gather(const DScene S, const WfBuffers B, const WfParams P)
{
  const uint32_t n = min(B.counters[QB_CNT_GATHER + P.depth], B.capacity);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 g0 = B.g0[i], g1 = B.g1[i], g2 = B.g2[i], g3 = B.g3[i], g4 = B.g4[i];
    const V3 p = v3(g0.x, g0.y, g0.z), N = v3(g0.w, g1.x, g1.y), V = v3(g1.z, g1.w, g2.x);
    const V3 Kd = v3(g2.y, g2.z, g2.w), Ks = v3(g3.x, g3.y, g3.z), T = v3(g4.x, g4.y, g4.z);
    const uint32_t w = __float_as_uint(g4.w);
    const int slot = (int) (w & 0x3FFFFFFFu);
    V3 c = v3(0.f);
    if (w & 0x40000000u) c = c + photon_gather_term(P.pm[0], p, N, V, Kd, Ks, g3.w);
    if (w & 0x80000000u) c = c + photon_gather_term(P.pm[1], p, N, V, Kd, Ks, g3.w);
    if (c.x != 0.f || c.y != 0.f || c.z != 0.f) accum_add(P.accum, slot, T * c);
  }
}
void launch_photon_gather(const DScene &S, const WfBuffers &B, const WfParams &P, cudaStream_t st)
{
  k_photon_gather<<<qb_sms() * 8, 128, 0, st>>>(S, B, P);
}

// C-ABI helper (qg_photon_estimate): EstimateIrradiance at caller-supplied points
__global__ void k_photon_estimate(const DPhotonMap M, const float *__restrict__ pos, const float *__restrict__ nrm, uint32_t n, float *__restrict__ irrad,
                                  float *__restrict__ dir)
{
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    V3 I, D;
    photon_estimate(M, v3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]), v3(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]), I, D);
    irrad[3 * i] = I.x; irrad[3 * i + 1] = I.y; irrad[3 * i + 2] = I.z;
    dir[3 * i] = D.x; dir[3 * i + 1] = D.y; dir[3 * i + 2] = D.z;
  }
}
void launch_photon_estimate(const DPhotonMap &M, const float *pos, const float *nrm, uint32_t n, float *irrad, float *dir, cudaStream_t st)
{
  k_photon_estimate<<<qb_sms() * 4, 128, 0, st>>>(M, pos, nrm, n, irrad, dir);
}

// ---- emission pass -----------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t photon_key(uint32_t e, int map, uint32_t seed) { return rng_key(e, 0x50484F54u + (uint32_t) map, seed); }

__device__ __forceinline__ void put_ray(qgRay &q, V3 p, V3 d)
{
  q.p[0] = p.x; q.p[1] = p.y; q.p[2] = p.z;
  q.dir[0] = d.x; q.dir[1] = d.y; q.dir[2] = d.z;
  q.has_diff = 0;  // DiffRay(p, dir): x and y are copies of the centre ray (src/core/ray.h:51-56)
}

// renderer.cpp:145-163 (photon map) / :216-233 (caustics map): pick a light, RandomPhoton(), Normalize(), power = intensity / #lights
__global__ void k_photon_emit(const DScene S, const PhotonBatch Bt, uint32_t e0, uint32_t E, int map, uint32_t seed, int n_plights,
                              const int32_t *__restrict__ plights)
{
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < E; i += gridDim.x * blockDim.x) {
    Rng rng = {photon_key(e0 + i, map, seed), 0u};
    int id = 0;
    if (n_plights > 1) {
      const float r = rng.next();
      // the photon map floors, the caustics map ceils (renderer.cpp:151 against :222): kept as it is
      const float x = map == 0 ? floorf(r * (float) n_plights) : ceilf(r * (float) n_plights);
      id = min((int) x, n_plights - 1);
    }
    const DLight &lt = S.lights[plights[id]];
    const V3 dir = uniform_sphere(rng);  // PointLight::RandomPhoton (lights.cpp:76-80)
    put_ray(Bt.rays[0][i], ld3(lt.position), normalize(dir));
    const V3 I = ld3(lt.intensity) * (1.f / (float) n_plights);  // GetPhotonIntensity * lightScale
    Bt.state[i] = make_float4(I.x, I.y, I.z, __uint_as_float(0u));
    Bt.rng_n[i] = rng.n;
    Bt.count[i] = 0u;
    Bt.path_of[0][i] = i;
  }
}

// MtlBlinn_PhotonMap::RandomPhotonBounce (:503-575). hit = what TraceNodeNormal left; returns false when the photon ends here.
__device__ bool photon_bounce(const DScene &S, const DMaterial &M, const qgHit &hit, V3 ray_dir, Rng &rng, V3 &c, V3 &new_dir)
{
  const float color_luma_threshold = 0.00001f;
  const bool front = hit.front != 0, has_tex = hit.has_texture != 0;
  const V3 uvw = v3(hit.uvw[0], hit.uvw[1], hit.uvw[2]);
  const V3 d0 = v3(hit.duvw[0], hit.duvw[1], hit.duvw[2]), d1 = v3(hit.duvw[3], hit.duvw[4], hit.duvw[5]);
  const V3 V = -ray_dir;
  const V3 N = v3(hit.N[0], hit.N[1], hit.N[2]);
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
  const V3 Kd = texcolor_sample(S, M.diffuse, has_tex, uvw, d0, d1);
  const V3 Ks = texcolor_sample(S, M.specular, has_tex, uvw, d0, d1);
  // RandomSelectMtl (:107-150)
  const float lumaT = luma(sT), lumaR = luma(sR), lumaD = luma(Kd);
  const float r = rng.next();
  const float coefT = lumaT, coefR = coefT + lumaR, coefD = coefR + lumaD, coefA = coefD + M.kill;
  const float rcp = 1.f / coefA;
  const float select = r * coefA;
  V3 dir = v3(0.f), bxdf = v3(0.f);
  float pdf = 1.f, scale;
  bool ok = false;
  if (select < coefT && lumaT > color_luma_threshold) {  // SampleTransmitBxDF (:152-178)
    scale = lumaT * rcp;
    dir = tDir;
    if (M.refr_gloss > 0.f)
      for (int it = 0; it < 64; ++it) {
        dir = normalize(normalize(tDir) + uniform_ball(rng, 2.f * M.refr_gloss));
        if (!(dot(dir, Y) > 0)) break;
      }
    bxdf = sT;
    ok = true;
  } else if (select < coefR && lumaR > color_luma_threshold) {  // SampleReflectionBxDF (:179-203)
    scale = lumaR * rcp;
    dir = rDir;
    if (M.refl_gloss > 0.f)
      for (int it = 0; it < 64; ++it) {
        dir = normalize(normalize(rDir) + uniform_ball(rng, 2.f * M.refl_gloss));
        if (!(dot(dir, Y) < 0)) break;
      }
    bxdf = sR;
    ok = true;
  } else if (select < coefD && lumaD > color_luma_threshold) {  // SampleDiffuseBxDF with photonMap = true (:204-224)
    scale = lumaD * rcp;
    if (front) {
      dir = to_local_frame(N, uniform_hemisphere(rng));
      const V3 L = normalize(dir);
      const V3 H = normalize(V + L);
      const float cosNH = qmax(0.f, dot(N, H));
      bxdf = Kd + Ks * powf(cosNH, M.gloss);
      pdf = 0.5f;
      ok = true;
    }
  } else {
    scale = coefA * rcp;  // ABSORB
  }
  if (!ok) return false;
  new_dir = normalize(dir);  // ray = DiffRay(p, sampleDir); ray.Normalize()
  c = c * bxdf / (pdf * scale);
  if (!front) c = c * v3(expf(-M.absorption[0] * hit.z), expf(-M.absorption[1] * hit.z), expf(-M.absorption[2] * hit.z));  // Attenuation
  return true;
}

// the body of the emission loops (renderer.cpp:164-192 / :234-265) for one traced segment of every photon still flying
__global__ void k_photon_step(const DScene S, const PhotonBatch Bt, int cur, uint32_t n_cur, int map, uint32_t e0, uint32_t seed)
{
  for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < n_cur; q += gridDim.x * blockDim.x) {
    const qgHit hit = Bt.hits[q];
    if (!hit.hit) continue;  // the photon left the scene
    const uint32_t path = Bt.path_of[cur][q];
    const qgRay &ray = Bt.rays[cur][q];
    const V3 ray_dir = v3(ray.dir[0], ray.dir[1], ray.dir[2]);
    const int mi = S.nodes[hit.node].material;
    if (mi < 0) continue;  // a node without material (the reference dereferences a null pointer here)
    const DMaterial &M = S.materials[mi];
    const bool multi = M.kind == QG_MTL_MULTI;
    // IsPhotonSurface(0): MtlBlinn: luma of the untextured diffuse colour > 0 (MtlBlinn_PhotonMap.h:74-77); MultiMtl
    // inherits Material's `true` — and its RandomPhotonBounce `false` (core/material.h:55-63)
    const bool surface = multi || luma(ld3(M.diffuse.color)) > 0.f;
    float4 st = Bt.state[path];
    uint32_t packed = __float_as_uint(st.w);
    const uint32_t bounce = packed & 255u;
    const bool diffuse_hit = (packed >> 8) & 1u;
    V3 c = v3(st.x, st.y, st.z);
    if (surface && bounce != 0u && (map == 0 || !diffuse_hit)) {
      const uint32_t k = Bt.count[path];
      if (k < (uint32_t) Bt.max_bounce) {
        float *o = Bt.rec + ((size_t) path * Bt.max_bounce + k) * 9;
        o[0] = hit.p[0]; o[1] = hit.p[1]; o[2] = hit.p[2];
        o[3] = ray_dir.x; o[4] = ray_dir.y; o[5] = ray_dir.z;
        o[6] = c.x; o[7] = c.y; o[8] = c.z;
        Bt.count[path] = k + 1u;
      }
    }
    if (multi) continue;
    Rng rng = {photon_key(e0 + path, map, seed), Bt.rng_n[path]};
    V3 nd;
    if (!photon_bounce(S, M, hit, ray_dir, rng, c, nd)) continue;
    const uint32_t nb = bounce + 1u;
    // caustics map: hInfo.c.hasDiffuseHit = diffuseHit || mtl->IsPhotonSurface(0) (renderer.cpp:259-262)
    packed = nb | ((diffuse_hit || surface) ? 256u : 0u);
    Bt.state[path] = make_float4(c.x, c.y, c.z, __uint_as_float(packed));
    Bt.rng_n[path] = rng.n;
    if (nb >= (uint32_t) Bt.max_bounce) continue;  // while (bounce < photonMapBounce)
    const uint32_t o = atomicAdd(Bt.n_next, 1u);
    put_ray(Bt.rays[cur ^ 1][o], v3(hit.p[0], hit.p[1], hit.p[2]), nd);
    Bt.path_of[cur ^ 1][o] = path;
  }
}

void launch_photon_emit(const DScene &S, const PhotonBatch &Bt, uint32_t e0, uint32_t E, int map, uint32_t seed, int n_plights, const int32_t *plights,
                        cudaStream_t st)
{
  k_photon_emit<<<qb_sms() * 4, 128, 0, st>>>(S, Bt, e0, E, map, seed, n_plights, plights);
}
void launch_photon_step(const DScene &S, const PhotonBatch &Bt, int cur, uint32_t n_cur, int map, uint32_t e0, uint32_t seed, cudaStream_t st)
{
  k_photon_step<<<qb_sms() * 4, 128, 0, st>>>(S, Bt, cur, n_cur, map, e0, seed);
}

}  // namespace qb
