// wave_rays.cuh — what the traversal kernels (trace_kernels.cu) and the fused shade kernel (shade_kernels.cu) both need of a
// wave's rays: the primary DiffRay of a (pixel, sample), and the compact hit record trace leaves for finish / shade.
// Both translation units are compiled with --fmad=false: the rays and the hit attributes must round like the reference's.
#pragma once
#include "traverse.cuh"
#include "wavefront.h"

namespace qb {

__device__ __forceinline__ float4 f4(float x, float y, float z, float w) { return make_float4(x, y, z, w); }

// primary DiffRay of pixel (i,j), sample s: texpos = (i + H11(s), j + H13(s)), origin = wf_campos (lens sample with DOF)
__device__ __forceinline__ void primary_ray(const WfParams &P, int i, int j, int s, Ray3 &c, Ray3 &x, Ray3 &y)
{
  const float tx = halton(s, 11) + (float) i;
  const float ty = halton(s, 13) + (float) j;
  const V3 cpt = P.scr_a + tx * P.scr_u + ty * P.scr_v;
  const V3 xpt = P.scr_a + (tx + 0.01f) * P.scr_u + ty * P.scr_v;
  const V3 ypt = P.scr_a + tx * P.scr_u + (ty + 0.01f) * P.scr_v;
  const V3 campos = wf_campos(P, i, j, s);
  c.p = x.p = y.p = campos;
  c.d = normalize(cpt - campos);
  x.d = normalize(xpt - campos);
  y.d = normalize(ypt - campos);
}

// compact hit record of the wavefront: r0 = z, k, slot, tex_slot ; r1 = tex_t, tex_k, last_tri, -
__device__ __forceinline__ HitRec unpack_rec(const float4 a, const float4 b)
{
  HitRec h;
  h.z = a.x; h.k = __float_as_int(a.y); h.slot = __float_as_int(a.z); h.tex_slot = __float_as_int(a.w);
  h.tex_t = b.x; h.tex_k = __float_as_int(b.y); h.last_tri = __float_as_int(b.z);
  return h;
}
__device__ __forceinline__ HitRec get_rec(const float4 *r0, const float4 *r1, uint32_t idx) { return unpack_rec(r0[idx], r1[idx]); }

}  // namespace qb
