// ref_math.h — float3 arithmetic of the reference render path, restated for the host with the
// SAME rounding sequence the reference's device code has after nvcc 12.9 contracts it.
//
// TEST INFRASTRUCTURE ONLY (part of the CPU oracle; never linked into libcrt_b200.so).
//
// The reference writes plain `a*b + c*d` expressions (reference: CudaRayTracer/src/Core/Vec3.cuh:205-234)
// and is built with nvcc defaults (-fmad=true). Which products get fused decides the low bits of every
// hit distance, so the oracle spells the fused form out with fmaf() and is compiled with
// -ffp-contract=off. The forms below were read off the PTX/SASS of the reference kernels built by
// oracle/Makefile (see DESIGN.md "FMA contraction table"):
//   dot(u,v)        = fma(u.z,v.z, fma(u.x,v.x, u.y*v.y))
//   cross(u,v).x    = fma(u.y,v.z, -(u.z*v.y))      (y, z cyclic)
//   lengthSquared   = fma(z,z, fma(x,x, y*y))
//   o + t*d         = fma(t,d,o)
//   v - k*n         = fma(-n,k,v)
//   1/x, sqrt, a/b  = IEEE round-to-nearest (rcp.rn / sqrt.rn / div.rn)
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

struct V3 {
    float x, y, z;
};

static inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
static inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
static inline V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
static inline V3 operator*(float t, V3 a) { return V3{t * a.x, t * a.y, t * a.z}; }
static inline V3 operator*(V3 a, float t) { return V3{a.x * t, a.y * t, a.z * t}; }

// reference: Vec3.cuh:214-216
static inline float dot(V3 u, V3 v) { return fmaf(u.z, v.z, fmaf(u.x, v.x, u.y * v.y)); }
// reference: Vec3.cuh:218-222
static inline V3 cross(V3 u, V3 v) {
    return V3{fmaf(u.y, v.z, -(u.z * v.y)), fmaf(u.z, v.x, -(u.x * v.z)), fmaf(u.x, v.y, -(u.y * v.x))};
}
// reference: Vec3.cuh:95-97
static inline float length_squared(V3 v) { return fmaf(v.z, v.z, fmaf(v.x, v.x, v.y * v.y)); }
// reference: Vec3.cuh:211-213 with operator/ = (1/t)*v (Vec3.cuh:195-197)
static inline V3 unit_vector(V3 v) {
    float inv = 1.0f / sqrtf(length_squared(v));
    return V3{v.x * inv, v.y * inv, v.z * inv};
}
// reference: Ray.cuh:14-17
static inline V3 point_at(V3 o, V3 d, float t) { return V3{fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z)}; }
// reference: Vec3.cuh:223-226; 2*dot is formed as dot+dot (exact), then fused multiply-subtract
static inline V3 reflect(V3 v, V3 n) {
    float d = dot(v, n);
    float k = d + d;
    return V3{fmaf(-n.x, k, v.x), fmaf(-n.y, k, v.y), fmaf(-n.z, k, v.z)};
}
// dot(-u, n) as the reference evaluates it inside Dielectric::scatter / refract. SASS of the sm_100 build:
// FMUL t = u.x*n.x; FFMA t = -u.y*n.y - t; FFMA -u.z*n.z + t   (x is the plain multiply here, unlike dot())
static inline float dot_neg(V3 u, V3 n) { return fmaf(-u.z, n.z, fmaf(-u.y, n.y, -(u.x * n.x))); }
// reference: Vec3.cuh:228-233
static inline V3 refract(V3 uv, V3 n, float eta) {
    float cos_theta = fminf(dot_neg(uv, n), 1.0f);
    V3 perp = V3{eta * fmaf(n.x, cos_theta, uv.x), eta * fmaf(n.y, cos_theta, uv.y), eta * fmaf(n.z, cos_theta, uv.z)};
    float s = sqrtf(fabsf(1.0f - length_squared(perp)));
    return V3{fmaf(-s, n.x, perp.x), fmaf(-s, n.y, perp.y), fmaf(-s, n.z, perp.z)};
}
// reference: Vec3.cuh:99-102
static inline bool near_zero(V3 v) {
    const float s = 1e-8f;
    return (fabsf(v.x) < s) && (fabsf(v.y) < s) && (fabsf(v.z) < s);
}
static inline float comp(const V3& v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

static inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

}  // namespace orc
