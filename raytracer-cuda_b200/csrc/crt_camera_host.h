// crt_camera_host.h — host-side camera set-up: crt_camera (pose + lens) -> derived vectors.
//
// Replaces Camera::updateCameraVectors, which the reference runs on the HOST before uploading the
// camera to constant memory (reference: CudaRayTracer/src/Core/Camera.cuh:159-182, uploaded by
// CUDARenderer::updateCamera, src/CUDARenderer.cuh:51-53). Plain float arithmetic in the reference's
// statement order; libm's cosf/sinf/tanf like the reference's host code. This file must be compiled
// without FMA contraction (x86-64 baseline has none; build flags add -ffp-contract=off).
#pragma once
#include <cmath>
#include "crt_types.h"

namespace crt {

inline vec3 host_unit(vec3 v) {
    float len = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    float inv = 1 / len;
    return mk3(inv * v.x, inv * v.y, inv * v.z);
}
inline vec3 host_cross(vec3 u, vec3 v) { return mk3(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x); }

inline Camera make_camera(const crt_camera& c) {
    const float PI = 3.1415926535897932385f;   // Core/Utility.cuh:9
    Camera k;
    k.pos = mk3(c.position[0], c.position[1], c.position[2]);
    vec3 front;
    front.x = -cosf(c.yaw_deg * PI / 180.0f) * cosf(c.pitch_deg * PI / 180.0f);
    front.y = -sinf(c.pitch_deg * PI / 180.0f);
    front.z = -sinf(c.yaw_deg * PI / 180.0f) * cosf(c.pitch_deg * PI / 180.0f);
    front = host_unit(front);
    const vec3 wup = mk3(c.world_up[0], c.world_up[1], c.world_up[2]);
    k.right = host_unit(host_cross(front, wup));
    k.up = host_unit(host_cross(k.right, front));
    const float theta = c.vfov_deg * PI / 180.0f;
    const float h = tanf(theta / 2.0f);
    const float vh = 2.0f * h;
    const float vw = c.aspect * vh;
    const float sh = c.focus_dist * vw, sv = c.focus_dist * vh;
    k.horiz = mk3(sh * k.right.x, sh * k.right.y, sh * k.right.z);
    k.vert = mk3(sv * k.up.x, sv * k.up.y, sv * k.up.z);
    const vec3 hh = mk3(0.5f * k.horiz.x, 0.5f * k.horiz.y, 0.5f * k.horiz.z);
    const vec3 hv = mk3(0.5f * k.vert.x, 0.5f * k.vert.y, 0.5f * k.vert.z);
    const vec3 ff = mk3(c.focus_dist * front.x, c.focus_dist * front.y, c.focus_dist * front.z);
    k.ll = mk3(((k.pos.x - hh.x) - hv.x) - ff.x, ((k.pos.y - hh.y) - hv.y) - ff.y, ((k.pos.z - hh.z) - hv.z) - ff.z);
    k.lens_radius = c.aperture / 2.0f;
    return k;
}

inline MatRec make_mat_record(const crt_material& m) {
    MatRec r;
    const bool light = m.type == CRT_MAT_DIFFUSE_LIGHT;
    r.a.x = light ? m.emission[0] : m.albedo[0];
    r.a.y = light ? m.emission[1] : m.albedo[1];
    r.a.z = light ? m.emission[2] : m.albedo[2];
    r.a.w = m.roughness;
    r.ior = m.ior;
    r.type = m.type;
    r.pad0 = r.pad1 = 0;
    return r;
}

}  // namespace crt
