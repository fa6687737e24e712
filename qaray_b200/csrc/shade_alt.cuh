// shade_alt.cuh — the other classes the reference's `MtlBlinn` alias can select (src/materials/materials.h:57-61),
// as wavefront shading steps. Included by shade_kernels.cu after its texture / sampling helpers.
//
//   QG_CLASS_PATHTRACING    MtlBlinn_PathTracing::Shade    src/materials/MtlBlinn_PathTracing.cpp:69-300   1 child / hit
//   QG_CLASS_MONTECARLOGI   MtlBlinn_MonteCarloGI::Shade   src/materials/MtlBlinn_MonteCarloGI.cpp:34-266  <= 2 + 10 children
//   QG_CLASS_BLINN_BASIC    MtlBlinn_Basic::Shade          src/materials/MtlBlinn_Basic.cpp:30-185         <= 2 children
//   QG_CLASS_PHONG_BASIC    MtlPhong_Basic::Shade          src/materials/MtlPhong_Basic.cpp:30-186         <= 2 children
//
// The reference recurses; here every child ray becomes its own path of the next depth level carrying the product of
// the factors its ancestors multiply its radiance by (throughput), so a level-d launch shades all hits of recursion
// depth d at once and every path adds `throughput * local terms` to its pixel. The last three classes trace their
// reflection / refraction rays WITH propagated differential rays (DiffRay(p, dir, px, dxdir, py, dydir)); those ride in
// ps4..ps6 and the differential hit distances come back in h4.
//
// Random numbers: a child path gets its own counter-based stream keyed by (parent key, parent counter, child index) —
// the reference's depth-first draw order cannot be reproduced breadth-first, and is not meaningful (time-seeded).
// MtlBlinn_PathTracing has one child, so it keeps one stream per sample, consumed in the reference's order.
#pragma once

namespace qb {

struct AltChild {
  bool on, diff;
  V3 dir, T, absorb;
  V3 xp, xd, yp, yd;
  uint32_t key, n, flags;
};

// warp-converged: every lane calls it; lanes with c.on append one path to the queue of depth d + 1
__device__ __forceinline__ void push_child(const WfBuffers &B, const WfParams &P, int d, int lane, uint32_t lt_mask, V3 p, int slot,
                                           const AltChild &c, OutRes &rp)
{
  // a child whose throughput is exactly zero can add nothing to the pixel: not traced (like zero-contribution shadow rays)
  const bool on = c.on && (c.T.x != 0.f || c.T.y != 0.f || c.T.z != 0.f);
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, on);
  if (!m) return;
  const int nb = (d & 1) ^ 1;
  const uint32_t q = res_take(rp, &B.counters[QB_CNT_PATHS + d + 1], lane, (uint32_t) __popc(m), (uint32_t) __popc(m & lt_mask));
  if (!on) return;
  if (q >= B.capacity) return;  // cannot happen: the wave size divides the capacity by the class's worst-case growth
  const V3 nd = normalize(c.dir);  // DiffRay::Normalize
  B.ps0[nb][q] = make_float4(p.x, p.y, p.z, nd.x);
  B.ps1[nb][q] = make_float4(nd.y, nd.z, __int_as_float(slot), __uint_as_float(c.key));
  B.ps2[nb][q] = make_float4(c.T.x, c.T.y, c.T.z, __uint_as_float((c.n << 8) | c.flags | (c.diff ? QB_FLAG_HAS_DIFF : 0u)));
  B.ps3[nb][q] = make_float4(c.absorb.x, c.absorb.y, c.absorb.z, 0.f);
  if (P.carry_diff && c.diff) {
    const V3 xd = normalize(c.xd), yd = normalize(c.yd);
    B.ps4[nb][q] = make_float4(c.xp.x, c.xp.y, c.xp.z, xd.x);
    B.ps5[nb][q] = make_float4(xd.y, xd.z, c.yp.x, c.yp.y);
    B.ps6[nb][q] = make_float4(c.yp.z, yd.x, yd.y, yd.z);
  }
}

// the factor a light's Illuminate() value is multiplied by, per class
template <int CLS>
__device__ __forceinline__ V3 light_factor(V3 N, V3 V, V3 L, V3 Kd, V3 Ks, float gloss, float norm_di)
{
  const float cosNL = qmax(0.f, dot(N, L));
  if (CLS == QG_CLASS_PHONG_BASIC) {  // MtlPhong_Basic.cpp:176-180
    const V3 R = 2.f * dot(L, N) * N - normalize(L);
    const float cosVR = qmax(0.f, dot(V, R));
    return Kd * cosNL + Ks * powf(cosVR, gloss);
  }
  const V3 H = normalize(V + L);
  const float cosNH = qmax(0.f, dot(N, H));
  if (CLS == QG_CLASS_PATHTRACING)  // MtlBlinn_PathTracing.cpp:160-172: normCoefDI enters twice
    return (norm_di * norm_di * cosNL) * (Kd + Ks * powf(cosNH, gloss));
  if (CLS == QG_CLASS_MONTECARLOGI)  // MtlBlinn_MonteCarloGI.cpp:184-194: no cosNL on the specular lobe
    return Kd * cosNL + Ks * powf(cosNH, gloss);
  return Kd * cosNL + Ks * (powf(cosNH, gloss) * cosNL);  // MtlBlinn_Basic.cpp:171-180
}

// PRIMARY = depth 0 (the ray is rebuilt from the pixel) / depth >= 1 (read from the path state): two instantiations,
// because nvcc 12.9 merges the two sources of the ray origin wrongly when they meet in one kernel (it kept the
// camera's x for both branches — the same defect k_wf_trace works around; found by the class parity test).
template <int CLS, bool PRIMARY>
__global__ void __launch_bounds__(QB_BLOCK) k_wf_shade_alt(const DScene S, const WfBuffers B, const WfParams P)
{
  constexpr bool WHITTED = CLS == QG_CLASS_BLINN_BASIC || CLS == QG_CLASS_PHONG_BASIC;
  constexpr bool BRANCHING = CLS != QG_CLASS_PATHTRACING;
  const int d = P.depth;
  const uint32_t n = PRIMARY ? P.n_paths : min(B.counters[QB_CNT_PATHS + d], B.capacity);
  const int bounce_count = P.max_bounce - d;
  const int lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int pb = d & 1;
  const float norm_di = S.num_lights > 0 ? 1.f / (float) S.num_lights : 1.f;
  // MtlBlinn_MonteCarloGI.cpp:176-178: 10 hemisphere samples at the camera-visible hit, 1 below
  const int num_mc = CLS == QG_CLASS_MONTECARLOGI ? ((P.max_bounce - 1 - bounce_count >= 0) ? 1 : 10) : 0;
  OutRes rp = {0u, 0u}, rs = {0u, 0u};  // reservations in the path queue of depth d + 1 and the shadow queue of depth d (k_wf_shade)
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

    bool shade = false, lights_on = false, front = false;
    V3 p = v3(0.f), N = v3(0.f), V = v3(0.f), T = v3(1.f), Kd = v3(0.f), Ks = v3(0.f);
    float gloss = 0.f;
    int slot = 0;
    Rng rng = {0u, 0u};
    AltChild c0, c1;  // PATHTRACING: c0 = the sampled lobe. Others: c0 = refraction, c1 = reflection
    c0.on = c1.on = false; c0.diff = c1.diff = false;
    c0.flags = c1.flags = 0u;
    c0.absorb = c1.absorb = v3(0.f);
    c0.dir = c1.dir = c0.T = c1.T = v3(0.f);
    c0.xp = c0.xd = c0.yp = c0.yd = c1.xp = c1.xd = c1.yp = c1.yd = v3(0.f);
    c0.key = c1.key = c0.n = c1.n = 0u;
    V3 mc_factor_d = v3(0.f), mc_factor_s = v3(0.f);  // MONTECARLOGI: Kd and Ks of the hit, for the hemisphere samples

    if (idx < n) {
      const float4 h0 = B.h0[idx];
      const int node = __float_as_int(h0.y);
      Ray3 rc, rx, ry;
      V3 absorb_prev = v3(0.f);
      bool valid = true;
      if (PRIMARY) {
        int i, j, s;
        valid = wf_decode(P, idx, i, j, s, slot);
        if (valid) {
          rng.key = rng_key((uint32_t) (j * P.width + i), (uint32_t) s, P.seed);
          rng.n = wf_dof_draws(P);
          atomicAdd(P.accum + 4ll * slot + 3, 1.0f);
          if (P.depth_plane && (s == 0 || P.mode == 1)) P.depth_plane[slot] = node >= 0 ? h0.x : QB_BIGFLOAT;
          const float tx = halton(s, 11) + (float) i, ty = halton(s, 13) + (float) j;
          if (node < 0) {  // background by screen position (renderer.cpp:336-338)
            accum_add(P.accum, slot, texcolor_sample1(S, S.background, v3(tx / (float) P.width, ty / (float) P.height, 0.f)));
          } else {  // the primary DiffRay (renderer.cpp:312-326)
            const V3 campos = wf_campos(P, i, j, s);
            rc.p = rx.p = ry.p = campos;
            rc.d = normalize(P.scr_a + tx * P.scr_u + ty * P.scr_v - campos);
            rx.d = normalize(P.scr_a + (tx + 0.01f) * P.scr_u + ty * P.scr_v - campos);
            ry.d = normalize(P.scr_a + tx * P.scr_u + (ty + 0.01f) * P.scr_v - campos);
          }
        }
      } else {
        const float4 a = B.ps0[pb][idx], b = B.ps1[pb][idx], c = B.ps2[pb][idx], e = B.ps3[pb][idx];
        rc.p = v3(a.x, a.y, a.z);
        rc.d = v3(a.w, b.x, b.y);
        rx = rc;
        ry = rc;
        slot = __float_as_int(b.z);
        rng.key = __float_as_uint(b.w);
        T = v3(c.x, c.y, c.z);
        const uint32_t packed = __float_as_uint(c.w);
        rng.n = packed >> 8;
        absorb_prev = v3(e.x, e.y, e.z);
        if (P.carry_diff && (packed & QB_FLAG_HAS_DIFF)) {
          const float4 f = B.ps4[pb][idx], g = B.ps5[pb][idx], h = B.ps6[pb][idx];
          rx.p = v3(f.x, f.y, f.z); rx.d = v3(f.w, g.x, g.y);
          ry.p = v3(g.z, g.w, h.x); ry.d = v3(h.y, h.z, h.w);
        }
        valid = slot >= 0;  // slot -1: a dead entry (unused tail of a shade warp's reservation)
        if (node < 0 && valid) accum_add(P.accum, slot, T * sample_environment(S, rc.d));
      }
      if (valid && node >= 0) {
        const float4 h1 = B.h1[idx], h2 = B.h2[idx], h3 = B.h3[idx];
        const float z = h0.x;
        const int mtl_id = __float_as_int(h0.z);
        const int hflags = __float_as_int(h0.w);
        front = hflags & 1;
        const bool has_tex = hflags & 2;
        p = v3(h1.x, h1.y, h1.z);
        N = normalize(v3(h1.w, h2.x, h2.y));
        const V3 uvw = v3(h2.z, h2.w, 0.f);
        const V3 d0 = v3(h3.x, h3.y, 0.f), d1 = v3(h3.z, h3.w, 0.f);
        if (!PRIMARY && !front) T = T * v3(expf(-absorb_prev.x * z), expf(-absorb_prev.y * z), expf(-absorb_prev.z * z));
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
          shade = true;
          V = normalize(-rc.d);
          const V3 Vx = normalize(-rx.d), Vy = normalize(-ry.d);
          V3 px = p, py = p;
          if (P.carry_diff) {
            const float4 h4 = B.h4[idx];
            px = rx.p + rx.d * h4.x;
            py = ry.p + ry.d * h4.y;
          }
#ifdef QB_DEBUG_PIX
          if (P.mode == 1) printf("d=%d idx=%u node=%d front=%d z=%g p=(%g %g %g) rc.p=(%g %g %g) rc.d=(%g %g %g) rx.p=(%g %g %g) rx.d=(%g %g %g) h4=(%g %g) px=(%g %g %g) d0=(%g %g) d1=(%g %g) T=(%g %g %g)\n",
            d, idx, node, (int) front, z, p.x, p.y, p.z, rc.p.x, rc.p.y, rc.p.z, rc.d.x, rc.d.y, rc.d.z, rx.p.x, rx.p.y, rx.p.z, rx.d.x, rx.d.y, rx.d.z,
            P.carry_diff ? B.h4[idx].x : 0.f, P.carry_diff ? B.h4[idx].y : 0.f, px.x, px.y, px.z, d0.x, d0.y, d1.x, d1.y, T.x, T.y, T.z);
#endif
          V3 color = v3(0.f);
          if (!WHITTED) color = texcolor_sample(S, M.emission, has_tex, uvw, d0, d1);
          V3 X, Y;
          if (WHITTED) {  // MtlBlinn_Basic.cpp:48-50
            X = normalize(cross(cross(N, V), N));
            Y = normalize(N * dot(N, V));
          } else {
            Y = dot(N, V) > 0.f ? N : -N;
            X = normalize(cross(Y, cross(V, Y)));
          }
          const float nIOR = front ? 1.f / M.ior : M.ior;
          const V3 tK0 = texcolor_sample(S, M.refraction, has_tex, uvw, d0, d1);
          const V3 rK0 = texcolor_sample(S, M.reflection, has_tex, uvw, d0, d1);
          Ks = texcolor_sample(S, M.specular, has_tex, uvw, d0, d1);
          Kd = texcolor_sample(S, M.diffuse, has_tex, uvw, d0, d1);
          gloss = M.gloss;
          if (CLS == QG_CLASS_PATHTRACING) {
            // ---- MtlBlinn_PathTracing.cpp:113-300 ------------------------------------------------------------------------
            const float cosI = dot(N, V);
            const float sinI = sqrtf(1 - cosI * cosI);
            const float sinO = qmax(0.f, qmin(1.f, sinI * nIOR));
            const float cosO = sqrtf(1.f - sinO * sinO);
            const V3 tDir = -X * sinO - Y * cosO;
            const V3 rDir = 2.f * N * (dot(N, V)) - V;
            const float C0 = (nIOR - 1.f) * (nIOR - 1.f) / ((nIOR + 1.f) * (nIOR + 1.f));
            const float rC = C0 + (1.f - C0) * powf(1.f - qabs(cosI), 5.f);
            const float tC = 1.f - rC;
            const bool tot = (nIOR * sinI) > 1.001f;
            const V3 sT = tot ? v3(0.f) : tK0 * tC;
            const V3 sR = tot ? (rK0 + tK0) : (rK0 + tK0 * rC);
            const float select = rng.next();
            float coefT = qmax(sT.x, qmax(sT.y, sT.z)), coefR = qmax(sR.x, qmax(sR.y, sR.z));
            float coefS = qmax(Ks.x, qmax(Ks.y, Ks.z)), coefD = qmax(Kd.x, qmax(Kd.y, Kd.z));
            const float coefSum = 0.f + coefT + coefR + coefS + coefD;
            coefT /= coefSum; coefR /= coefSum; coefS /= coefSum; coefD /= coefSum;
            const float sumT = coefT + 0.f, sumR = coefR + coefT + 0.f, sumS = coefS + coefR + coefT + 0.f;
            lights_on = true;
            if (bounce_count > 0) {
              const V3 nZ = Y;
              const V3 nY = (qabs(nZ.x) > qabs(nZ.y)) ? normalize(v3(nZ.z, 0, -nZ.x)) : normalize(v3(0, -nZ.z, nZ.y));
              const V3 nX = normalize(cross(nY, nZ));
              // SetReflectionGlossiness / SetRefractionGlossiness (:57-67): 1/gloss above 0.001, else -1
              const float g_t = M.refr_gloss > 0.001f ? 1.f / M.refr_gloss : -1.f;
              const float g_r = M.refl_gloss > 0.001f ? 1.f / M.refl_gloss : -1.f;
              V3 dir = v3(0.f), BxDF = v3(0.f);
              float PDF = 1.f;
              bool go = false;
              if (select <= sumT && coefT > 1e-6f) {
                if (g_t > 0.f) {
                  const V3 sm = normalize(cos_weighted_hemisphere(rng));
                  dir = -(sm.x * nX + sm.y * nY + sm.z * nZ);
                  BxDF = sT * powf(qmax(0.f, dot(V, tDir)), g_t);
                } else {
                  dir = tDir;
                  BxDF = sT;
                }
                PDF = coefT;
                go = true;
              } else if (select < sumR && coefR > 1e-6f) {
                if (g_r > 0.f) {
                  const V3 sm = normalize(cos_weighted_hemisphere(rng));
                  dir = sm.x * nX + sm.y * nY + sm.z * nZ;
                  BxDF = sR * powf(qmax(0.f, dot(V, rDir)), g_r);
                } else {
                  dir = rDir;
                  BxDF = sR;
                }
                PDF = coefR;
                go = true;
              } else if (select < sumS && coefS > 1e-6f) {
                if (gloss > 0.f && front) {
                  const V3 sm = normalize(cos_weighted_hemisphere(rng));
                  dir = sm.x * nX + sm.y * nY + sm.z * nZ;
                  const V3 H = normalize(V + normalize(dir));
                  BxDF = Ks * powf(qmax(0.f, dot(N, H)), gloss);
                  PDF = coefS;
                  go = true;
                }
              } else if (coefD > 1e-6f) {
                if (front) {
                  const V3 sm = normalize(cos_weighted_hemisphere(rng));
                  dir = sm.x * nX + sm.y * nY + sm.z * nZ;
                  BxDF = Kd;
                  PDF = coefD;
                  go = true;
                }
              }
              if (go) {
                c0.on = true;
                c0.dir = dir;
                c0.T = T * (BxDF / PDF);
                c0.key = rng.key;  // one child: the sample keeps its stream
              }
            }
          } else {
            // ---- the Whitted-style classes: MtlBlinn_Basic.cpp:56-163, MtlBlinn_MonteCarloGI.cpp:63-165 --------------------
            const float th_gloss = 0.001f;
            const float th_color = WHITTED ? 0.01f : 0.001f;
            const float th_total = WHITTED ? 1.01f : 1.001f;
            V3 tjN = N, rjN = N;
            if (M.refr_gloss > th_gloss) tjN = normalize(N + uniform_ball(rng, M.refr_gloss));
            if (M.refl_gloss > th_gloss) rjN = normalize(N + uniform_ball(rng, M.refl_gloss));
            const float cosI = dot(tjN, V);
            const float sinI = sqrtf(1 - cosI * cosI);
            const float cosIx = dot(tjN, Vx), sinIx = sqrtf(1 - cosIx * cosIx);
            const float cosIy = dot(tjN, Vy), sinIy = sqrtf(1 - cosIy * cosIy);
            const float sinO = qmax(0.f, qmin(1.f, sinI * nIOR)), cosO = sqrtf(1.f - sinO * sinO);
            const float sinOx = qmax(0.f, qmin(1.f, sinIx * nIOR)), cosOx = sqrtf(1.f - sinOx * sinOx);
            const float sinOy = qmax(0.f, qmin(1.f, sinIy * nIOR)), cosOy = sqrtf(1.f - sinOy * sinOy);
            const V3 tDir = -X * sinO - Y * cosO, txDir = -X * sinOx - Y * cosOx, tyDir = -X * sinOy - Y * cosOy;
            const V3 rDir = 2.f * rjN * (dot(rjN, V)) - V, rxDir = 2.f * rjN * (dot(rjN, Vx)) - Vx, ryDir = 2.f * rjN * (dot(rjN, Vy)) - Vy;
            // (the reference's do-while around this block repeats only when a direction leaves its half space, with
            //  unchanged inputs when nothing is glossy — it never iterates twice in a terminating run)
            const float C0 = (nIOR - 1.f) * (nIOR - 1.f) / ((nIOR + 1.f) * (nIOR + 1.f));
            const float rC = C0 + (1.f - C0) * powf(1.f - qabs(cosI), 5.f);
            const float tC = 1.f - rC;
            const bool tot = (nIOR * sinI) > th_total;
            const V3 tK = tot ? v3(0.f) : tK0 * tC;
            const V3 rK = tot ? (rK0 + tK0) : (rK0 + tK0 * rC);
            const V3 absorb = ld3(M.absorption);
            if (bounce_count > 0 && (tK.x > th_color || tK.y > th_color || tK.z > th_color)) {
              c0.on = true; c0.diff = true;
              c0.dir = tDir; c0.xp = px; c0.xd = txDir; c0.yp = py; c0.yd = tyDir;
              c0.T = T * tK; c0.absorb = absorb;
              c0.key = child_key(rng.key, rng.n, 0u);
            }
            if (bounce_count > 0 && (rK.x > th_color || rK.y > th_color || rK.z > th_color)) {
              c1.on = true; c1.diff = true;
              c1.dir = rDir; c1.xp = px; c1.xd = rxDir; c1.yp = py; c1.yd = ryDir;
              c1.T = T * rK; c1.absorb = absorb;
              c1.key = child_key(rng.key, rng.n, 1u);
            }
            lights_on = front;  // both classes light front faces only (MtlBlinn_Basic.cpp:164, MtlBlinn_MonteCarloGI.cpp:179)
            if (CLS == QG_CLASS_MONTECARLOGI && front && bounce_count > 0) {
              mc_factor_d = Kd;
              mc_factor_s = Ks;
            }
          }
          if (color.x != 0.f || color.y != 0.f || color.z != 0.f) accum_add(P.accum, slot, T * color);
        }
      }
    }

    // ---- lights: ambient adds Kd * I directly (Whitted classes only, MtlBlinn_Basic.cpp:166-168); every other light
    // queues one shadow ray carrying the contribution it unlocks (GenLight::Shadow, lights.cpp:39-48)
    for (int l = 0; l < S.num_lights; ++l) {
      const DLight &lt = S.lights[l];
      if (lt.type == QG_LIGHT_AMBIENT) {  // uniform across the warp
        if (WHITTED && lights_on) {
          const V3 c = T * (Kd * ld3(lt.intensity));
          if (c.x != 0.f || c.y != 0.f || c.z != 0.f) accum_add(P.accum, slot, c);
        }
        continue;
      }
      const bool area = lt.type != QG_LIGHT_DIRECT && lt.size > 0.01f;  // lights.cpp:52,87 (uniform across the warp)
      bool push = false;
      V3 sdir = v3(0.f), contrib = v3(0.f);
      float tmax = QB_BIGFLOAT;
      if (lights_on) {
        V3 I = ld3(lt.intensity), Ldir;
        if (lt.type == QG_LIGHT_DIRECT) {
          sdir = normalize(-ld3(lt.direction));
          Ldir = ld3(lt.direction);
        } else {
          const V3 dv = ld3(lt.position) - p;
          sdir = normalize(dv);
          tmax = length(dv);
          if (!P.no_falloff && !area) I = I * qmin(1.f, 1.f / length2(dv));  // InverseSquareFalloff, lights.cpp:23-30
          Ldir = normalize(p - ld3(lt.position));
          if (lt.type == QG_LIGHT_SPOT) I = I * spot_attenuation(lt, Ldir);
        }
        const V3 L = normalize(-Ldir);
        contrib = T * (I * light_factor<CLS>(N, V, L, Kd, Ks, gloss, norm_di));
        push = (contrib.x != 0.f) || (contrib.y != 0.f) || (contrib.z != 0.f);
      }
      if (area) {
        push_area_group(B, d, lane, lt_mask, push, p, l, slot, contrib, child_key(rng.key, rng.n, 0x100u + (uint32_t) l));
        continue;
      }
      const uint32_t mask = __ballot_sync(0xFFFFFFFFu, push);
      if (mask) {
        const uint32_t q = res_take(rs, &B.counters[QB_CNT_SHADOW + d], lane, (uint32_t) __popc(mask), (uint32_t) __popc(mask & lt_mask));
        if (push) {
          if (q < B.shadow_capacity) {
            B.s0[q] = make_float4(p.x, p.y, p.z, sdir.x);
            B.s1[q] = make_float4(sdir.y, sdir.z, tmax, __int_as_float(slot));
            B.s2[q] = make_float4(contrib.x, contrib.y, contrib.z, 0.f);
          }
        }
      }
    }

    // ---- children ----------------------------------------------------------------------------------------------------
    if (!BRANCHING) c0.n = rng.n;
    push_child(B, P, d, lane, lt_mask, p, slot, c0, rp);
    if (BRANCHING) push_child(B, P, d, lane, lt_mask, p, slot, c1, rp);
    if (CLS == QG_CLASS_MONTECARLOGI) {
      // MtlBlinn_MonteCarloGI.cpp:197-257: cos-weighted hemisphere samples about N, each weighted
      // (cosNL * Ks * cosNH^gloss + Kd) / numSampleMC; traced as DiffRay(p, dir) (no differentials), no absorption
      const bool mc = shade && front && (P.max_bounce - d) > 0;
      const float norm_gi = 1.f / (float) num_mc;
      for (int i = 0; i < num_mc; ++i) {
        AltChild cm;
        cm.on = false; cm.diff = false; cm.flags = 0u; cm.n = 0u; cm.key = 0u;
        cm.absorb = cm.xp = cm.xd = cm.yp = cm.yd = cm.dir = cm.T = v3(0.f);
        if (mc) {
          Rng r2 = {child_key(rng.key, rng.n, 2u + (uint32_t) i), 0u};
          const V3 sm = normalize(cos_weighted_hemisphere(r2));
          const V3 nz = N;
          const V3 ny = (qabs(nz.x) > qabs(nz.y)) ? normalize(v3(nz.z, 0, -nz.x)) : normalize(v3(0, -nz.z, nz.y));
          const V3 nx = normalize(cross(ny, nz));
          const V3 dirMC = sm.x * nx + sm.y * ny + sm.z * nz;
          const V3 H = normalize(V + dirMC);
          const float cosNL = qmax(0.f, dot(N, dirMC)), cosNH = qmax(0.f, dot(N, H));
          cm.on = true;
          cm.dir = dirMC;
          cm.T = T * ((cosNL * mc_factor_s * powf(cosNH, gloss) + mc_factor_d) * norm_gi);
          cm.key = r2.key;
          cm.n = r2.n;
        }
        push_child(B, P, d, lane, lt_mask, p, slot, cm, rp);
      }
    }
  }
  res_flush(B, d, lane, rp, rs);
}

}  // namespace qb
