// crt_shade.h — RNG stream, camera rays, tagged-union BSDF, background, output packing.
//
// Replaces the virtual Material hierarchy and the cuRAND-state-by-pointer helpers
// (reference: CudaRayTracer/src/Core/Material.cuh:49-150, Core/Utility.cuh:18-76,
// Core/Camera.cuh:32-44, CRTUtility.cuh:9-38). Draw order per sample is the reference's
// (SURVEY.md §3.3-13): lens disk (x,y per try) -> u jitter -> v jitter -> per bounce
// [RR draw if bounce >= min] -> scatter draws.
#pragma once
#include "crt_trace.h"

namespace crt {

// XORWOW (Marsaglia 2003) as cuRAND steps it: curand() = xorwow + Weyl counter d += 362437.
// The state is seeded by cuRAND's own curand_init in the rng-init kernel; only the step is here.
struct Rng {
    uint32_t v0, v1, v2, v3, v4, d;
};

CRT_HD uint32_t rng_next(Rng& r) {
    uint32_t t = r.v0 ^ (r.v0 >> 2);
    r.v0 = r.v1; r.v1 = r.v2; r.v2 = r.v3; r.v3 = r.v4;
    r.v4 = (r.v4 ^ (r.v4 << 4)) ^ (t ^ (t << 1));
    r.d += 362437u;
    return r.v4 + r.d;
}
// curand_uniform: fma(x, 2^-32, 2^-33), in (0,1]
CRT_HD float rng_uniform(Rng& r) { return f_fma((float)rng_next(r), 2.3283064365386963e-10f, 1.1641532182693481e-10f); }
// Utility::randomFloat(-1,1) = fma(u, 2, -1)                                (Utility.cuh:18-21)
CRT_HD float rng_pm1(Rng& r) { return f_fma(rng_uniform(r), 2.0f, -1.0f); }

// Utility::randomPointInUnitSphere (Utility.cuh:45-53): x, y, z drawn in that order
CRT_HD vec3 rng_in_unit_sphere(Rng& r) {
    while (true) {
        float x = rng_pm1(r);
        float y = rng_pm1(r);
        float z = rng_pm1(r);
        vec3 p = mk3(x, y, z);
        if (v_len2(p) >= 1.f) continue;
        return p;
    }
}

// Camera::getRay (Camera.cuh:32-44). Direction is NOT normalised.
CRT_HD void camera_ray(const Camera& k, int px, int py, int W, int H, Rng& rng, vec3& o, vec3& d) {
    float dx, dy;
    while (true) {   // Utility::randomPointInUnitDisk (Utility.cuh:55-62)
        dx = rng_pm1(rng);
        dy = rng_pm1(rng);
        if (f_fma(dx, dx, f_mul(dy, dy)) >= 1.f) continue;
        break;
    }
    const float rdx = f_mul(k.lens_radius, dx), rdy = f_mul(k.lens_radius, dy);
    const vec3 off = mk3(f_fma(rdx, k.right.x, f_mul(rdy, k.up.x)), f_fma(rdx, k.right.y, f_mul(rdy, k.up.y)),
                         f_fma(rdx, k.right.z, f_mul(rdy, k.up.z)));
    const float u = f_div(f_add(rng_uniform(rng), (float)px), (float)W);
    const float v = f_div(f_add(rng_uniform(rng), (float)py), (float)H);
    o = v_add(k.pos, off);
    const vec3 t = mk3(f_fma(v, k.vert.x, f_fma(u, k.horiz.x, k.ll.x)), f_fma(v, k.vert.y, f_fma(u, k.horiz.y, k.ll.y)),
                       f_fma(v, k.vert.z, f_fma(u, k.horiz.z, k.ll.z)));
    d = v_sub(v_sub(t, k.pos), off);
}

// getBackgroundColor (CRTUtility.cuh:34-38)
CRT_HD vec3 background(vec3 d) {
    const float inv = f_rcp(f_sqrt(v_len2(d)));
    const float t = f_mul(f_fma(inv, d.y, 1.0f), 0.5f);
    const float omt = f_sub(1.0f, t);
    return mk3(f_fma(t, 0.5f, omt), f_fma(t, 0.7f, omt), f_add(t, omt));
}

struct Surface {
    vec3 point, normal;
    bool front;
    uint32_t mat;
};

// Hit point, geometric normal flipped against the ray, material
// (Mesh.cuh:295-305, Sphere.cuh:38-44, HitInfo.cuh:15-18)
CRT_HD Surface surface_at(const SceneView& sc, const HitRec& h, vec3 o, vec3 d) {
    Surface s;
    const f4 q0 = ld_f4(&sc.prims[h.prim].q0);
    const f4 q1 = ld_f4(&sc.prims[h.prim].q1);
    const f4 q2 = ld_f4(&sc.prims[h.prim].q2);
    s.point = v_at(o, d, h.t);
    s.mat = f_bits(q2.z);
    vec3 n;
    if (f_bits(q2.w) == (uint32_t)PRIM_TRIANGLE) {
        n = v_unit(v_cross(mk3(q0.w, q1.x, q1.y), mk3(q1.z, q1.w, q2.x)));
    } else {
        n = v_scale(v_sub(s.point, mk3(q0.x, q0.y, q0.z)), f_rcp(q0.w));
    }
    s.front = v_dot(d, n) < 0.f;
    s.normal = s.front ? n : v_neg(n);
    return s;
}

// scatter(): returns true when the path continues (att, o, d updated); false = emit() is the
// path's radiance (Material.cuh:51-55, CUDAKernels.h:128-134).
CRT_HD bool scatter(const MatRec& m, const Surface& s, vec3 in_d, Rng& rng, vec3& att, vec3& o, vec3& d) {
    if (m.type == CRT_MAT_LAMBERTIAN) {                       // Material.cuh:66-77
        const vec3 p = rng_in_unit_sphere(rng);
        const float inv = f_rcp(f_sqrt(v_len2(p)));
        vec3 dir = mk3(f_fma(p.x, inv, s.normal.x), f_fma(p.y, inv, s.normal.y), f_fma(p.z, inv, s.normal.z));
        if (v_near_zero(dir)) dir = s.normal;
        o = s.point;
        d = dir;
        att = mk3(m.a.x, m.a.y, m.a.z);
        return true;
    }
    if (m.type == CRT_MAT_METAL) {                            // Material.cuh:86-96
        const float rough = m.a.w < 1.f ? m.a.w : 1.f;
        const vec3 refl = v_unit(v_reflect(in_d, s.normal));
        const vec3 p = rng_in_unit_sphere(rng);
        const vec3 u = v_scale(p, f_rcp(f_sqrt(v_len2(p))));
        const vec3 dir = mk3(f_fma(rough, u.x, refl.x), f_fma(rough, u.y, refl.y), f_fma(rough, u.z, refl.z));
        o = s.point;
        d = dir;
        att = mk3(m.a.x, m.a.y, m.a.z);
        return v_dot(dir, s.normal) > 0.f;
    }
    if (m.type == CRT_MAT_DIELECTRIC) {                       // Material.cuh:109-137
        att = mk3(1.f, 1.f, 1.f);
        const float ri = s.front ? f_rcp(m.ior) : m.ior;
        const vec3 ud = v_unit(in_d);
        const float cosf32 = fminf(v_dot_neg(ud, s.normal), 1.0f);
        const double cos_theta = (double)cosf32;
        const double sin_theta = sqrt(1.0 - cos_theta * cos_theta);
        bool do_reflect = (double)ri * sin_theta > 1.0;
        if (!do_reflect) {
            // schlicksReflectance: pow(float,int) resolves to double pow; one double fma; rounded to float
            float r0 = f_div(f_sub(1.f, ri), f_add(1.f, ri));
            r0 = f_mul(r0, r0);
            const float one_minus = f_sub(1.f, cosf32);
            const double pw = pow((double)one_minus, 5.0);
            const float refl = (float)fma(pw, (double)f_sub(1.f, r0), (double)r0);
            const float u = rng_uniform(rng);
            do_reflect = refl > u;
        }
        o = s.point;
        d = do_reflect ? v_reflect(ud, s.normal) : v_refract(ud, s.normal, ri);
        return true;
    }
    return false;                                             // DiffuseLight
}

CRT_HD vec3 emitted(const MatRec& m) {                        // Material.cuh:53-55,144-146
    if (m.type == CRT_MAT_DIFFUSE_LIGHT) return mk3(m.a.x, m.a.y, m.a.z);
    return mk3(0.f, 0.f, 0.f);
}

// writeColor (CRTUtility.cuh:21-32): gamma 2 through the double overload, clamp to 0.999, *256
CRT_HD uint32_t pack_channel(float c) {
    float g = c > 0.f ? (float)sqrt((double)c) : 0.f;
    if (g > 0.999f) g = 0.999f;
    if (g < 0.f) g = 0.f;
    return (uint32_t)(f_mul(256.f, g)) & 0xffu;
}
CRT_HD uint32_t pack_rgba8(vec3 c) {
    return pack_channel(c.x) | (pack_channel(c.y) << 8) | (pack_channel(c.z) << 16) | 0xff000000u;
}

// One path: rayColor (CUDAKernels.h:102-145). Returns the sample's radiance; `rays` counts
// closest-hit queries.
template <bool COUNT>
CRT_HD vec3 path_radiance(const SceneView& sc, const RenderConst& rc, vec3 o, vec3 d, Rng& rng, uint32_t& rays, TraceCounters* ctr) {
    vec3 thr = mk3(1.f, 1.f, 1.f);
    for (int bounce = 0; bounce < rc.max_bounces; bounce++) {
        if (bounce >= rc.min_bounces) {
            const float p = fminf(fmaxf(thr.x, fmaxf(thr.y, thr.z)), rc.rr_max_prob);
            if (rng_uniform(rng) > p) break;
            thr = v_scale(thr, f_rcp(p));
        }
        HitRec h;
        rays++;
        if (trace_closest<COUNT>(sc, o, d, 0.001f, INFINITY, h, ctr)) {
            const Surface s = surface_at(sc, h, o, d);
            if (s.mat < sc.num_mats) {
                const MatRec m = sc.mats[s.mat];
                vec3 att, no, nd;
                if (scatter(m, s, d, rng, att, no, nd)) {
                    thr = v_mul(thr, att);
                    o = no;
                    d = nd;
                } else {
                    return emitted(m);
                }
            }
        } else {
            return v_mul(thr, background(d));
        }
    }
    return mk3(0.f, 0.f, 0.f);
}

}  // namespace crt
