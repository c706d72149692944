// crt_math.h — float3 arithmetic with an explicit rounding sequence.
//
// The render path has to agree bit-for-bit with the reference's compiled device code on hit
// distances (primary-hit ids are required to be bit-exact, BASELINE.json north_star). The reference
// writes `a*b + c*d` and lets nvcc contract it; this library instead spells every multiply/add with
// round-to-nearest intrinsics (__fmaf_rn / __fmul_rn / __fadd_rn ...), which the compiler never
// re-associates or fuses, in exactly the fused form the reference's SASS has
// (DESIGN.md "FMA contraction table"; reference: CudaRayTracer/src/Core/Vec3.cuh:205-234).
//
// CRT_HD code also compiles as plain C++ (g++) for the host-side simulation used by tests
// (tests/hostsim); the product library only ever runs the device instantiation.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define CRT_HD __host__ __device__ __forceinline__
#define CRT_D __device__ __forceinline__
#else
#define CRT_HD inline
#define CRT_D inline
#endif

namespace crt {

#if defined(__CUDA_ARCH__)
CRT_HD float f_fma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
CRT_HD float f_mul(float a, float b) { return __fmul_rn(a, b); }
CRT_HD float f_add(float a, float b) { return __fadd_rn(a, b); }
CRT_HD float f_sub(float a, float b) { return __fsub_rn(a, b); }
CRT_HD float f_rcp(float a) { return __frcp_rn(a); }
CRT_HD float f_sqrt(float a) { return __fsqrt_rn(a); }
CRT_HD float f_div(float a, float b) { return __fdiv_rn(a, b); }
CRT_HD uint32_t f_bits(float f) { return __float_as_uint(f); }
CRT_HD float bits_f(uint32_t u) { return __uint_as_float(u); }
#else
// host build: compile with -ffp-contract=off so these stay separate roundings
CRT_HD float f_fma(float a, float b, float c) { return fmaf(a, b, c); }
CRT_HD float f_mul(float a, float b) { return a * b; }
CRT_HD float f_add(float a, float b) { return a + b; }
CRT_HD float f_sub(float a, float b) { return a - b; }
CRT_HD float f_rcp(float a) { return 1.0f / a; }
CRT_HD float f_sqrt(float a) { return sqrtf(a); }
CRT_HD float f_div(float a, float b) { return a / b; }
CRT_HD uint32_t f_bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
CRT_HD float bits_f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
#endif

struct vec3 {
    float x, y, z;
};

CRT_HD vec3 mk3(float x, float y, float z) { vec3 v; v.x = x; v.y = y; v.z = z; return v; }
CRT_HD vec3 v_add(vec3 a, vec3 b) { return mk3(f_add(a.x, b.x), f_add(a.y, b.y), f_add(a.z, b.z)); }
CRT_HD vec3 v_sub(vec3 a, vec3 b) { return mk3(f_sub(a.x, b.x), f_sub(a.y, b.y), f_sub(a.z, b.z)); }
CRT_HD vec3 v_neg(vec3 a) { return mk3(-a.x, -a.y, -a.z); }
CRT_HD vec3 v_mul(vec3 a, vec3 b) { return mk3(f_mul(a.x, b.x), f_mul(a.y, b.y), f_mul(a.z, b.z)); }
CRT_HD vec3 v_scale(vec3 a, float t) { return mk3(f_mul(a.x, t), f_mul(a.y, t), f_mul(a.z, t)); }

// dot(u,v) = fma(u.z,v.z, fma(u.x,v.x, u.y*v.y))                       (Vec3.cuh:214-216)
CRT_HD float v_dot(vec3 u, vec3 v) { return f_fma(u.z, v.z, f_fma(u.x, v.x, f_mul(u.y, v.y))); }
// cross(u,v).x = fma(u.y,v.z, -(u.z*v.y)), cyclic                       (Vec3.cuh:218-222)
CRT_HD vec3 v_cross(vec3 u, vec3 v) {
    return mk3(f_fma(u.y, v.z, -f_mul(u.z, v.y)), f_fma(u.z, v.x, -f_mul(u.x, v.z)), f_fma(u.x, v.y, -f_mul(u.y, v.x)));
}
// lengthSquared = fma(z,z, fma(x,x, y*y))                               (Vec3.cuh:95-97)
CRT_HD float v_len2(vec3 v) { return f_fma(v.z, v.z, f_fma(v.x, v.x, f_mul(v.y, v.y))); }
// unitVector(v) = (1/sqrt(len2)) * v                                    (Vec3.cuh:195-197,211-213)
CRT_HD vec3 v_unit(vec3 v) { return v_scale(v, f_rcp(f_sqrt(v_len2(v)))); }
// o + t*d = fma(t,d,o)                                                  (Ray.cuh:14-17)
CRT_HD vec3 v_at(vec3 o, vec3 d, float t) { return mk3(f_fma(t, d.x, o.x), f_fma(t, d.y, o.y), f_fma(t, d.z, o.z)); }
// reflect(v,n) = v - (2 dot(v,n)) n, 2*dot formed as dot+dot, fused multiply-subtract (Vec3.cuh:223-226)
CRT_HD vec3 v_reflect(vec3 v, vec3 n) {
    float d = v_dot(v, n);
    float k = f_add(d, d);
    return mk3(f_fma(-n.x, k, v.x), f_fma(-n.y, k, v.y), f_fma(-n.z, k, v.z));
}
// dot(-u, n) as compiled inside Dielectric::scatter / refract: the reference's SASS multiplies the x
// terms plainly and fuses y then z (FMUL x*nx; FFMA -y*ny - t; FFMA -z*nz + t)
CRT_HD float v_dot_neg(vec3 u, vec3 n) { return f_fma(-u.z, n.z, f_fma(-u.y, n.y, -f_mul(u.x, n.x))); }
// refract(uv, n, eta)                                                   (Vec3.cuh:228-233)
CRT_HD vec3 v_refract(vec3 uv, vec3 n, float eta) {
    float c = fminf(v_dot_neg(uv, n), 1.0f);
    vec3 perp = mk3(f_mul(eta, f_fma(n.x, c, uv.x)), f_mul(eta, f_fma(n.y, c, uv.y)), f_mul(eta, f_fma(n.z, c, uv.z)));
    float s = f_sqrt(fabsf(f_sub(1.0f, v_len2(perp))));
    return mk3(f_fma(-s, n.x, perp.x), f_fma(-s, n.y, perp.y), f_fma(-s, n.z, perp.z));
}
// Vec3::nearZero                                                        (Vec3.cuh:99-102)
CRT_HD bool v_near_zero(vec3 v) {
    const float s = 1e-8f;
    return (fabsf(v.x) < s) && (fabsf(v.y) < s) && (fabsf(v.z) < s);
}

}  // namespace crt
