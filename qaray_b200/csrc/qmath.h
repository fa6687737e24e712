// qmath.h — fp32 vector / matrix helpers shared by the host loader and the CUDA kernels.
//
// Every expression is written in the evaluation order of the reference's math layer (glm 0.9.8.4, vendored by
// the reference under external/glm, used through src/math/math.h:64-100) so that, compiled without FMA
// contraction (host: -ffp-contract=off, device: --fmad=false), results are bit-identical to the reference:
//   dot(vec3)      tmp = a*b; tmp.x + tmp.y + tmp.z            glm/detail/func_geometric.inl compute_dot<tvec3>
//   cross          (x.y*y.z - y.y*x.z, ...)                    compute_cross
//   normalize      v * (1 / sqrt(dot(v,v)))                    compute_normalize + inversesqrt
//   mat3 * vec3    m[0][r]*v.x + m[1][r]*v.y + m[2][r]*v.z     glm/detail/type_mat3x3.inl operator*
//   inverse(mat3)  cofactors * (1/det)                         glm/detail/func_matrix.inl compute_inverse<tmat3x3>
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define QB_HD __host__ __device__ __forceinline__
#else
#define QB_HD inline
#endif

namespace qb {

struct V3 {
  float x, y, z;
};

QB_HD V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
QB_HD V3 v3(float s) { return v3(s, s, s); }
QB_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
QB_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
QB_HD V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
QB_HD V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
QB_HD V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
QB_HD V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
QB_HD V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
QB_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
QB_HD V3 &operator+=(V3 &a, V3 b) { a = a + b; return a; }
QB_HD V3 &operator*=(V3 &a, V3 b) { a = a * b; return a; }
QB_HD V3 &operator*=(V3 &a, float s) { a = a * s; return a; }
QB_HD float comp(const V3 &a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

QB_HD float dot(V3 a, V3 b)
{
  const V3 t = a * b;
  return t.x + t.y + t.z;
}
QB_HD V3 cross(V3 x, V3 y)
{
  return v3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
QB_HD float length2(V3 a) { return dot(a, a); }
QB_HD float length(V3 a) { return sqrtf(dot(a, a)); }
QB_HD V3 normalize(V3 a) { return a * (1.0f / sqrtf(dot(a, a))); }

// The reference's MIN/MAX/ABS macros (src/math/math.h:104-107), including their NaN behaviour.
QB_HD float qmin(float x, float y) { return x < y ? x : y; }
QB_HD float qmax(float x, float y) { return x > y ? x : y; }
QB_HD float qabs(float x) { return x > 0 ? x : -x; }

// TriangleArea (src/objects/objects.cpp:30-41): signed projected area (x2) with `axis` dropped, and the axis
// IntersectTriangle drops: the one where abs(N) is largest (objects.cpp:235-247)
QB_HD float tri_area(int axis, V3 A, V3 B, V3 C)
{
  if (axis == 0) return (B.y - A.y) * (C.z - A.z) - (C.y - A.y) * (B.z - A.z);
  if (axis == 1) return (B.x - A.x) * (C.z - A.z) - (C.x - A.x) * (B.z - A.z);
  return (B.x - A.x) * (C.y - A.y) - (C.x - A.x) * (B.y - A.y);
}
QB_HD int tri_axis(V3 N)
{
  const float ax = qabs(N.x), ay = qabs(N.y), az = qabs(N.z);
  if (ax > ay && ax > az) return 0;
  if (ay > az) return 1;
  return 2;
}

// column-major 3x3: c[col][row], like glm::mat3
struct M3 {
  float c[3][3];
};

QB_HD M3 m3_identity()
{
  M3 m;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) m.c[i][j] = i == j ? 1.0f : 0.0f;
  return m;
}
QB_HD V3 mul(const M3 &m, V3 v)
{
  return v3(m.c[0][0] * v.x + m.c[1][0] * v.y + m.c[2][0] * v.z,
            m.c[0][1] * v.x + m.c[1][1] * v.y + m.c[2][1] * v.z,
            m.c[0][2] * v.x + m.c[1][2] * v.y + m.c[2][2] * v.z);
}
// dot(column k, v) for each k — Transformation::TransposeMult (src/core/transform.cpp:49-56)
QB_HD V3 transpose_mul(const M3 &m, V3 v)
{
  return v3(dot(v3(m.c[0][0], m.c[0][1], m.c[0][2]), v),
            dot(v3(m.c[1][0], m.c[1][1], m.c[1][2]), v),
            dot(v3(m.c[2][0], m.c[2][1], m.c[2][2]), v));
}
QB_HD M3 mul(const M3 &a, const M3 &b)
{
  M3 r;
  for (int j = 0; j < 3; ++j)
    for (int i = 0; i < 3; ++i) r.c[j][i] = a.c[0][i] * b.c[j][0] + a.c[1][i] * b.c[j][1] + a.c[2][i] * b.c[j][2];
  return r;
}
QB_HD M3 inverse(const M3 &m)
{
  const float ood = 1.0f / (+m.c[0][0] * (m.c[1][1] * m.c[2][2] - m.c[2][1] * m.c[1][2])
                            - m.c[1][0] * (m.c[0][1] * m.c[2][2] - m.c[2][1] * m.c[0][2])
                            + m.c[2][0] * (m.c[0][1] * m.c[1][2] - m.c[1][1] * m.c[0][2]));
  M3 r;
  r.c[0][0] = +(m.c[1][1] * m.c[2][2] - m.c[2][1] * m.c[1][2]) * ood;
  r.c[1][0] = -(m.c[1][0] * m.c[2][2] - m.c[2][0] * m.c[1][2]) * ood;
  r.c[2][0] = +(m.c[1][0] * m.c[2][1] - m.c[2][0] * m.c[1][1]) * ood;
  r.c[0][1] = -(m.c[0][1] * m.c[2][2] - m.c[2][1] * m.c[0][2]) * ood;
  r.c[1][1] = +(m.c[0][0] * m.c[2][2] - m.c[2][0] * m.c[0][2]) * ood;
  r.c[2][1] = -(m.c[0][0] * m.c[2][1] - m.c[2][0] * m.c[0][1]) * ood;
  r.c[0][2] = +(m.c[0][1] * m.c[1][2] - m.c[1][1] * m.c[0][2]) * ood;
  r.c[1][2] = -(m.c[0][0] * m.c[1][2] - m.c[1][0] * m.c[0][2]) * ood;
  r.c[2][2] = +(m.c[0][0] * m.c[1][1] - m.c[1][0] * m.c[0][1]) * ood;
  return r;
}

// The reference's Transformation (src/core/transform.h:36-79): p' = itm*(p-pos); back: tm*p+pos; normals itm^T.
struct Xform {
  M3 tm, itm;
  V3 pos;
};
QB_HD V3 xf_to(const Xform &t, V3 p) { return mul(t.itm, p - t.pos); }
QB_HD V3 xf_from(const Xform &t, V3 p) { return mul(t.tm, p) + t.pos; }
QB_HD V3 xf_vec_from(const Xform &t, V3 n) { return transpose_mul(t.itm, n); }

// Halton radical inverse (src/core/sampler.cpp:31-40)
QB_HD float halton(int index, int base)
{
  float r = 0;
  float f = 1.0f / (float) base;
  for (int i = index; i > 0; i /= base) {
    r += f * (float) (i % base);
    f /= (float) base;
  }
  return r;
}

QB_HD float luma(V3 c) { return 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z; }

// Counter-based per-path random stream. key identifies (pixel, sample, seed); draw n is a pure function of
// (key, n), so a path's numbers do not depend on scheduling, batch size or GPU count. Mapped to [0,1] exactly
// as the reference's xorshift32 output is: float(u32) / 2^32 (src/samplers/Sampler_Marsaglia.cpp:43-53).
QB_HD uint32_t rng_mix(uint32_t h)
{
  h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
  return h;
}
QB_HD uint32_t rng_key(uint32_t pixel, uint32_t sample, uint32_t seed)
{
  return rng_mix(rng_mix(pixel * 0x9E3779B1u + 0x7F4A7C15u) ^ (sample * 0x85EBCA77u) ^ (seed * 0xC2B2AE3Du + 0x27D4EB2Fu));
}
QB_HD uint32_t rng_bits(uint32_t key, uint32_t n) { return rng_mix(key ^ rng_mix(n * 0x9E3779B9u + 0x632BE5ABu)); }
QB_HD float rng_float(uint32_t key, uint32_t n) { return (float) rng_bits(key, n) * 2.3283064365386963e-10f; }

}  // namespace qb
