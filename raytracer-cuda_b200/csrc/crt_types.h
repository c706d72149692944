// crt_types.h — device data layout of the render path (flat SoA scene, wide BVH, camera).
//
// Replaces the reference's device-side polymorphic scene graph
// (reference: CudaRayTracer/src/Core/{Hittable,HittableList,BVHNode,Mesh,Sphere,Material}.cuh)
// with flat arrays. See DESIGN.md "Data layout in HBM".
#pragma once
#include "crt_math.h"
#include "../../include/crt_api.h"

namespace crt {

// ---- primitive record: 48 B = 3 x 16-byte loads, stored in BVH leaf order -----------------
//   triangle: q0 = (v0.x, v0.y, v0.z, e1.x)  q1 = (e1.y, e1.z, e2.x, e2.y)  q2 = (e2.z, id, mat, kind=0)
//   sphere  : q0 = (c.x,  c.y,  c.z,  r)     q1 = (r*r, 0, 0, 0)            q2 = (0,    id, mat, kind=1)
// e1 = v1 - v0 and e2 = v2 - v0 are the SAME fp32 subtractions the reference performs per test
// (reference: Core/Mesh.cuh:277-278), so storing them changes no bits. id = caller's primitive id
// (triangle index, or num_triangles + sphere index).
struct f4 {
    float x, y, z, w;
};
struct PrimRec {
    f4 q0, q1, q2;
};
enum { PRIM_TRIANGLE = 0, PRIM_SPHERE = 1 };

// ---- material: tagged union, 32 B (reference: Core/Material.cuh:16-47 MaterialData, 36 B) -----
//   a = (albedo.rgb | emission.rgb, roughness)   b = (ior, type, 0, 0)
struct MatRec {
    f4 a;
    float ior;
    int32_t type;
    int32_t pad0, pad1;
};

// ---- compressed 8-wide BVH node, 80 B = 5 x 16-byte loads (after Ylitie, Karras, Laine 2017) ----
//   n0 = (p.x, p.y, p.z, [ex | ey<<8 | ez<<16 | imask<<24])
//   n1 = (child_base, prim_base, meta[0..3], meta[4..7])
//   n2 = (qlo_x[0..3], qlo_x[4..7], qlo_y[0..3], qlo_y[4..7])
//   n3 = (qlo_z[0..3], qlo_z[4..7], qhi_x[0..3], qhi_x[4..7])
//   n4 = (qhi_y[0..3], qhi_y[4..7], qhi_z[0..3], qhi_z[4..7])
// child i box = p + 2^(e-127) * [qlo, qhi]; imask bit i = child i is an inner node;
// meta[i]: 0 = empty; inner: (1<<5) | (24+i); leaf: (unary prim count << 5) | first prim offset.
struct u4 {
    uint32_t x, y, z, w;
};
struct WideNode {
    u4 n0, n1, n2, n3, n4;
};

struct Camera {   // derived vectors (reference: Core/Camera.cuh:188-208)
    vec3 pos, right, up, ll, horiz, vert;
    float lens_radius;
};

struct SceneView {   // device pointers
    const WideNode* nodes;
    const PrimRec* prims;
    const MatRec* mats;
    uint32_t num_mats;
    uint32_t num_nodes;
    uint32_t num_prims;
    uint32_t giant_base;    // prims[giant_base .. giant_base+giant_count) are outside the BVH: every ray tests them
    uint32_t giant_count;
    // reference-exact mode: a triangle's record carries its tie key (position in the reference's permuted order) in the
    // id field; key_to_id[key] is the caller's triangle id (null: key = id). Spheres: key = id = num_tris + index.
    const uint32_t* key_to_id;
    uint32_t num_tris;
};

struct RenderConst {
    int width, height;
    int spp;
    int max_bounces;
    int min_bounces;
    float rr_max_prob;
};

}  // namespace crt
