// oracle.cpp — CPU restatement of the reference render path.  TEST INFRASTRUCTURE ONLY.
//
// See oracle.h for who may use this. Everything here follows the reference's algorithm and
// evaluation order; every function cites the reference lines it restates
// (paths relative to /root/reference/CudaRayTracer/src). Arithmetic uses ref_math.h so that the
// rounding sequence equals the reference's compiled device code; build with -ffp-contract=off.
#include "oracle.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

#include "ref_math.h"
#include "xorwow_host.h"

using namespace orc;

namespace {

const float kInf = std::numeric_limits<float>::infinity();

// ---------------------------------------------------------------------------------------------
// Interval / AABB                                   reference: Core/Interval.cuh, Core/AABB.cuh
// ---------------------------------------------------------------------------------------------
struct Interval {
    float min = kInf, max = -kInf;   // Interval.cuh:7-9 (default = empty)
};

struct Box {
    Interval x, y, z;
};

// AABB::padToMinimums, AABB.cuh:181-186 (Interval::expand pads by delta/2 each side, Interval.cuh:42-45)
void pad_to_minimums(Box& b) {
    const float delta = 0.000001f;
    const float pad = delta / 2.f;
    if (b.x.max - b.x.min < delta) { b.x.min = b.x.min - pad; b.x.max = b.x.max + pad; }
    if (b.y.max - b.y.min < delta) { b.y.min = b.y.min - pad; b.y.max = b.y.max + pad; }
    if (b.z.max - b.z.min < delta) { b.z.min = b.z.min - pad; b.z.max = b.z.max + pad; }
}

// AABB_EMPTY = AABB(INTERVAL_EMPTY x3): the 3-interval ctor pads (AABB.cuh:16-20,187); on an
// empty interval size = -inf < delta, and inf -/+ pad stays +/-inf.
Box box_empty() {
    Box b;
    pad_to_minimums(b);
    return b;
}

// AABB::expand(point), AABB.cuh:51-59 (no padding)
void box_expand(Box& b, V3 p) {
    b.x.min = fminf(b.x.min, p.x); b.x.max = fmaxf(b.x.max, p.x);
    b.y.min = fminf(b.y.min, p.y); b.y.max = fmaxf(b.y.max, p.y);
    b.z.min = fminf(b.z.min, p.z); b.z.max = fmaxf(b.z.max, p.z);
}

// AABB::expand(const AABB&), AABB.cuh:82-88 (pads)
void box_expand_box(Box& b, const Box& o) {
    b.x.min = fminf(b.x.min, o.x.min); b.x.max = fmaxf(b.x.max, o.x.max);
    b.y.min = fminf(b.y.min, o.y.min); b.y.max = fmaxf(b.y.max, o.y.max);
    b.z.min = fminf(b.z.min, o.z.min); b.z.max = fmaxf(b.z.max, o.z.max);
    pad_to_minimums(b);
}

// AABB::combine, AABB.cuh:90-97 (3-interval ctor => pads)
Box box_combine(const Box& a, const Box& c) {
    Box b;
    b.x.min = fminf(a.x.min, c.x.min); b.x.max = fmaxf(a.x.max, c.x.max);
    b.y.min = fminf(a.y.min, c.y.min); b.y.max = fmaxf(a.y.max, c.y.max);
    b.z.min = fminf(a.z.min, c.z.min); b.z.max = fmaxf(a.z.max, c.z.max);
    pad_to_minimums(b);
    return b;
}

// AABB(const Point3& a, const Point3& b), AABB.cuh:29-36 (pads)
Box box_from_points(V3 a, V3 c) {
    Box b;
    b.x.min = fminf(a.x, c.x); b.x.max = fmaxf(a.x, c.x);
    b.y.min = fminf(a.y, c.y); b.y.max = fmaxf(a.y, c.y);
    b.z.min = fminf(a.z, c.z); b.z.max = fmaxf(a.z, c.z);
    pad_to_minimums(b);
    return b;
}

// AABB::area, AABB.cuh:73-80: 2*(ex*ey + ey*ez + ez*ex), contracted as
// fma(ez,ex, fma(ex,ey, ey*ez)) by the same rule as dot().
float box_area(const Box& b) {
    float ex = b.x.max - b.x.min, ey = b.y.max - b.y.min, ez = b.z.max - b.z.min;
    return 2.0f * fmaf(ez, ex, fmaf(ex, ey, ey * ez));
}

// AABB::longestAxis, AABB.cuh:148-154
int box_longest_axis(const Box& b) {
    float sx = b.x.max - b.x.min, sy = b.y.max - b.y.min, sz = b.z.max - b.z.min;
    if (sx > sy) return sx > sz ? 0 : 2;
    return sy > sz ? 1 : 2;
}

// AABB::hit, AABB.cuh:123-146. invD = 1/dir per call; miss iff tmax <= tmin.
bool box_hit(const Box& b, V3 o, V3 d, float tmin_in, float tmax_in) {
    float ix = 1.0f / d.x, iy = 1.0f / d.y, iz = 1.0f / d.z;
    float t0x = (b.x.min - o.x) * ix, t0y = (b.y.min - o.y) * iy, t0z = (b.z.min - o.z) * iz;
    float t1x = (b.x.max - o.x) * ix, t1y = (b.y.max - o.y) * iy, t1z = (b.z.max - o.z) * iz;
    float tmin = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
    float tmax = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    tmin = fmaxf(tmin, tmin_in);
    tmax = fminf(tmax, tmax_in);
    return !(tmax <= tmin);
}

// ---------------------------------------------------------------------------------------------
// scene structures
// ---------------------------------------------------------------------------------------------
struct BNode {                 // BVHNode fields, Core/BVHNode.cuh:158-164
    Box box;
    int left = 0, right = 0;
    int obj_index = 0, obj_count = 0;
    bool is_leaf = false;
};

struct Mesh {                  // Core/Mesh.cuh:112-119
    std::vector<V3> verts;            // m_Vertices (positions only; normals/uvs are never read)
    std::vector<uint32_t> indices;    // m_Indices, permuted in place by the builder
    std::vector<int> face_mat;        // m_FaceMaterialIds, permuted alongside
    std::vector<int> orig_id;         // oracle-only: original triangle id per permuted face
    std::vector<BNode> nodes;         // m_MeshBVH
    Box bbox;                         // m_BoundingBox (over ALL vertices, Mesh.cuh:39-47)
    int mat_offset = 0;               // m_MatIDOffset
    int first_tri = 0;                // global id of this mesh's first triangle
};

struct Sphere {                // Core/Sphere.cuh:49-52
    V3 c;
    float r, r2;
    uint32_t mat;
    Box bbox;
};

struct Object {                // a slot of HittableList::m_Objects
    int kind;                  // 0 = mesh, 1 = sphere
    int index;
};

}  // namespace

struct orc_scene {
    std::vector<Mesh> meshes;
    std::vector<Sphere> spheres;
    std::vector<Object> objects;           // meshes first, then spheres (CUDAKernels.h:64-73)
    std::vector<BNode> top;                // top-level BVH (BVHNode.cuh:21-84)
    std::vector<crt_material> materials;   // HittableList::m_Materials
    int num_triangles = 0;
};

namespace {

// ---------------------------------------------------------------------------------------------
// Mesh::buildBVHMesh and helpers                                   reference: Core/Mesh.cuh:121-264
// ---------------------------------------------------------------------------------------------
V3 tri_centroid(const Mesh& m, int i) {   // Mesh.cuh:251-256: (p0+p1+p2) * (1.f/3.f)
    V3 p0 = m.verts[m.indices[i]], p1 = m.verts[m.indices[i + 1]], p2 = m.verts[m.indices[i + 2]];
    const float third = 1.f / 3.f;
    V3 s = (p0 + p1) + p2;
    return V3{s.x * third, s.y * third, s.z * third};
}

void expand_tri(const Mesh& m, Box& b, int i) {   // Mesh.cuh:242-249
    box_expand(b, m.verts[m.indices[i]]);
    box_expand(b, m.verts[m.indices[i + 1]]);
    box_expand(b, m.verts[m.indices[i + 2]]);
}

float evaluate_sah(const Mesh& m, int axis, float pos, int start, int end) {   // Mesh.cuh:222-240
    Box lb = box_empty(), rb = box_empty();
    int lc = 0, rc = 0;
    for (int i = start; i < end; i += 3) {
        V3 c = tri_centroid(m, i);
        if (comp(c, axis) < pos) { lc++; expand_tri(m, lb, i); }
        else { rc++; expand_tri(m, rb, i); }
    }
    // leftCount * leftBox.area() + rightCount * rightBox.area(): the reference's SASS (initMesh, sm_100)
    // is FMUL(la, lc) then FFMA(ra, rc, .)  -> fma(rc, ra, lc*la); pinned by tests/golden/c2s_perm_i32.bin
    float cost = fmaf((float)rc, box_area(rb), (float)lc * box_area(lb));
    return cost < 1e-8f ? 1e-8f : cost;
}

void build_mesh_bvh(Mesh& m) {   // Mesh.cuh:121-219
    const int index_count = (int)m.indices.size();
    const int num_tris = index_count / 3;
    const int max_nodes = 2 * num_tris - 1;
    if (max_nodes <= 0) return;
    m.nodes.assign((size_t)max_nodes, BNode());
    struct Entry { int start, end, node; };
    std::vector<Entry> stack;   // reference: fixed 64 entries, unchecked (Mesh.cuh:128)
    int next = 0;
    BNode& root = m.nodes[next++];
    root.box = m.bbox;
    root.obj_index = 0;
    root.obj_count = index_count;
    root.is_leaf = false;
    stack.push_back({0, index_count, 0});
    while (!stack.empty()) {
        Entry cur = stack.back();
        stack.pop_back();
        const int start = cur.start, end = cur.end;
        BNode& node = m.nodes[cur.node];
        const int span = end - start;
        if (span <= 30 || (next + 1) >= max_nodes) {   // Mesh.cuh:148: leaf at <= 10 triangles
            node.is_leaf = true;
            node.obj_index = start;
            node.obj_count = span;
            Box b = box_empty();                         // computeTrianglesAABB, Mesh.cuh:258-264
            for (int i = start; i < start + span; i += 3) expand_tri(m, b, i);
            node.box = b;
            continue;
        }
        int best_axis = 0;
        float best_pos = 0.f, best_cost = 1e30f;
        for (int axis = 0; axis < 3; axis++) {           // Mesh.cuh:164-179
            float mn = 1e30f, mx = -1e30f;
            for (int i = start; i < end; i += 3) {
                float c = comp(tri_centroid(m, i), axis);
                if (c < mn) mn = c;
                if (c > mx) mx = c;
            }
            float mid = 0.5f * (mn + mx);
            float cost = evaluate_sah(m, axis, mid, start, end);
            if (cost < best_cost) { best_cost = cost; best_axis = axis; best_pos = mid; }
        }
        int mid = start;                                  // Lomuto partition, Mesh.cuh:182-198
        for (int i = start; i < end; i += 3) {
            float c = comp(tri_centroid(m, i), best_axis);
            if (c < best_pos) {
                std::swap(m.indices[mid], m.indices[i]);
                std::swap(m.indices[mid + 1], m.indices[i + 1]);
                std::swap(m.indices[mid + 2], m.indices[i + 2]);
                std::swap(m.face_mat[mid / 3], m.face_mat[i / 3]);
                std::swap(m.orig_id[mid / 3], m.orig_id[i / 3]);
                mid += 3;
            }
        }
        node.left = next++;
        node.right = next++;
        node.is_leaf = false;
        stack.push_back({mid, end, node.right});          // right pushed first, left popped first
        stack.push_back({start, mid, node.left});
    }
    for (int i = next - 1; i >= 0; i--) {                 // bottom-up refit, Mesh.cuh:211-218
        BNode& nd = m.nodes[i];
        if (!nd.is_leaf) nd.box = box_combine(m.nodes[nd.left].box, m.nodes[nd.right].box);
    }
    m.nodes.resize((size_t)next);
}

const Box& object_box(const orc_scene& s, const Object& o) {
    return o.kind == 0 ? s.meshes[(size_t)o.index].bbox : s.spheres[(size_t)o.index].bbox;
}

// BVHNode::buildBVHScene, Core/BVHNode.cuh:21-84: median split in insertion order (the sort calls
// are commented out, :64-67), one object per leaf, node ids handed out by pre-increment.
void build_top_bvh(orc_scene& s) {
    const int n = (int)s.objects.size();
    if (n <= 0) return;
    s.top.assign((size_t)(2 * n - 1), BNode());
    struct Entry { int start, end, node; };
    std::vector<Entry> stack;
    int next = 0;
    stack.push_back({0, n, 0});
    while (!stack.empty()) {
        Entry cur = stack.back();
        stack.pop_back();
        BNode& node = s.top[(size_t)cur.node];
        node = BNode();
        node.box = box_empty();
        for (int i = cur.start; i < cur.end; i++) box_expand_box(node.box, object_box(s, s.objects[(size_t)i]));
        int span = cur.end - cur.start;
        if (span == 1) {
            node.obj_index = cur.start;
            node.obj_count = 1;
            node.is_leaf = true;
        } else {
            int mid = cur.start + span / 2;
            int l = ++next, r = ++next;
            node.left = l;
            node.right = r;
            node.is_leaf = false;
            stack.push_back({mid, cur.end, r});
            stack.push_back({cur.start, mid, l});
        }
    }
}

}  // namespace

namespace {

// ---------------------------------------------------------------------------------------------
// hit record + primitive tests
// ---------------------------------------------------------------------------------------------
struct Hit {                   // Core/HitInfo.cuh:7-19
    V3 point, normal;
    uint32_t mat = 0;
    float t = 0.f;
    bool front = false;
    int prim = -1;             // oracle-only: original primitive id
};

// HitInfo::setFaceNormal, HitInfo.cuh:15-18
void set_face_normal(Hit& h, V3 dir, V3 outward) {
    h.front = dot(dir, outward) < 0.f;
    h.normal = h.front ? outward : -outward;
}

// Mesh::rayTriangleIntersect, Mesh.cuh:266-308 (Moeller-Trumbore, eps 1e-8, inclusive edges)
bool hit_triangle(const Mesh& m, int i, V3 o, V3 d, float tmin, float tmax, Hit& rec) {
    const float EPSILON = 1e-8f;
    V3 p0 = m.verts[m.indices[i]], p1 = m.verts[m.indices[i + 1]], p2 = m.verts[m.indices[i + 2]];
    V3 e1 = p1 - p0, e2 = p2 - p0;
    V3 h = cross(d, e2);
    float a = dot(e1, h);
    if (fabsf(a) < EPSILON) return false;
    float f = 1.f / a;
    V3 s = o - p0;
    float u = f * dot(s, h);
    if (u < 0.f || u > 1.f) return false;
    V3 q = cross(s, e1);
    float v = f * dot(d, q);
    if (v < 0.f || (u + v) > 1.f) return false;
    float t = f * dot(e2, q);
    if (t < tmin || t > tmax) return false;
    rec.t = t;
    rec.point = point_at(o, d, t);
    int face = i / 3;
    rec.mat = (uint32_t)(m.face_mat[(size_t)face] + m.mat_offset);
    rec.prim = m.orig_id[(size_t)face];
    V3 n = unit_vector(cross(e1, e2));
    set_face_normal(rec, d, n);
    return true;
}

// Sphere::hit, Core/Sphere.cuh:18-47
bool hit_sphere(const Sphere& sp, int prim_id, V3 o, V3 d, float tmin, float tmax, Hit& rec) {
    V3 oc = o - sp.c;
    float a = dot(d, d);
    float half_b = dot(oc, d);
    float c = dot(oc, oc) - sp.r2;
    float disc = fmaf(half_b, half_b, -(a * c));       // half_b*half_b - a*c (FFMA with FMUL'd a*c)
    if (disc < 0) return false;
    float sq = sqrtf(disc);
    float root = (-half_b - sq) / a;
    if (root < tmin || root > tmax) {                   // Interval::outOfInterval, Interval.cuh:33-35
        root = (-half_b + sq) / a;                      // compiled as (sq - half_b) / a: same value
        if (root < tmin || root > tmax) return false;
    }
    rec.t = root;
    rec.point = point_at(o, d, root);
    rec.mat = sp.mat;
    rec.prim = prim_id;
    float inv_r = 1.0f / sp.r;                          // (P - C) / r = (1/r) * (P - C)
    V3 pc = rec.point - sp.c;
    V3 n = V3{pc.x * inv_r, pc.y * inv_r, pc.z * inv_r};
    set_face_normal(rec, d, n);
    return true;
}

// Mesh::hit, Mesh.cuh:55-110: left-first DFS, box test against [tmin, closestSoFar]
bool hit_mesh(const Mesh& m, V3 o, V3 d, float tmin, float tmax, Hit& rec, int mode) {
    bool any = false;
    float closest = tmax;
    if (mode == 1) {   // brute force in the reference's visiting order (permuted index order)
        for (int i = 0; i < (int)m.indices.size(); i += 3) {
            Hit tmp;
            if (hit_triangle(m, i, o, d, tmin, closest, tmp)) { any = true; closest = tmp.t; rec = tmp; }
        }
        return any;
    }
    if (m.nodes.empty()) return false;
    int stack[256];
    int sp = 0, idx = 0;
    while (true) {
        const BNode& node = m.nodes[(size_t)idx];
        if (box_hit(node.box, o, d, tmin, closest)) {
            if (!node.is_leaf) {
                stack[sp++] = node.right;
                idx = node.left;
                continue;
            }
            for (int i = node.obj_index; i < node.obj_index + node.obj_count; i += 3) {
                Hit tmp;
                if (hit_triangle(m, i, o, d, tmin, closest, tmp)) { any = true; closest = tmp.t; rec = tmp; }
            }
        }
        if (sp == 0) break;
        idx = stack[--sp];
    }
    return any;
}

bool hit_object(const orc_scene& s, int obj, V3 o, V3 d, float tmin, float tmax, Hit& rec, int mode) {
    const Object& ob = s.objects[(size_t)obj];
    if (ob.kind == 0) return hit_mesh(s.meshes[(size_t)ob.index], o, d, tmin, tmax, rec, mode);
    return hit_sphere(s.spheres[(size_t)ob.index], s.num_triangles + ob.index, o, d, tmin, tmax, rec);
}

// BVHNode::hit, Core/BVHNode.cuh:115-156: node boxes tested against the UNSHRUNK interval,
// objects against [tmin, closestSoFar].
bool hit_scene(const orc_scene& s, V3 o, V3 d, float tmin, float tmax, Hit& rec, int mode) {
    bool any = false;
    float closest = tmax;
    if (mode == 1) {
        for (int i = 0; i < (int)s.objects.size(); i++) {
            Hit tmp;
            if (hit_object(s, i, o, d, tmin, closest, tmp, mode)) { any = true; closest = tmp.t; rec = tmp; }
        }
        return any;
    }
    if (s.top.empty()) return false;
    int stack[1024];
    int sp = 0, idx = 0;
    while (true) {
        const BNode& node = s.top[(size_t)idx];
        if (box_hit(node.box, o, d, tmin, tmax)) {
            if (node.is_leaf) {
                Hit tmp;
                if (hit_object(s, node.obj_index, o, d, tmin, closest, tmp, mode)) { any = true; closest = tmp.t; rec = tmp; }
                if (sp == 0) break;
                idx = stack[--sp];
            } else {
                stack[sp++] = node.right;
                idx = node.left;
            }
        } else {
            if (sp == 0) break;
            idx = stack[--sp];
        }
    }
    return any;
}

// ---------------------------------------------------------------------------------------------
// RNG helpers                                                       reference: Core/Utility.cuh
// ---------------------------------------------------------------------------------------------
// randomFloat(-1, 1) = min + (max-min)*u, compiled to fma(u, 2, -1)          Utility.cuh:18-21
inline float rand_pm1(Xorwow& r) { return fmaf(xorwow_uniform(r), 2.0f, -1.0f); }

// randomPointInUnitSphere, Utility.cuh:45-53: 3 draws per try, x first (nvcc evaluates the
// Vec3(rand,rand,rand) arguments left to right), reject when |p|^2 >= 1
V3 random_in_unit_sphere(Xorwow& r) {
    while (true) {
        float x = rand_pm1(r), y = rand_pm1(r), z = rand_pm1(r);
        V3 p{x, y, z};
        if (length_squared(p) >= 1.f) continue;
        return p;
    }
}

// ---------------------------------------------------------------------------------------------
// materials                                                         reference: Core/Material.cuh
// ---------------------------------------------------------------------------------------------
// returns true when the ray scatters (attenuation + new ray set)
bool scatter(const crt_material& m, V3 in_o, V3 in_d, const Hit& rec, V3& att, V3& out_o, V3& out_d, Xorwow& rng) {
    (void)in_o;
    switch (m.type) {
    case CRT_MAT_LAMBERTIAN: {   // Material.cuh:66-77
        V3 p = random_in_unit_sphere(rng);
        float inv = 1.0f / sqrtf(length_squared(p));
        // rec.Normal + unitVector(p): the scale by 1/len and the add fuse into one fma per axis
        V3 dir{fmaf(p.x, inv, rec.normal.x), fmaf(p.y, inv, rec.normal.y), fmaf(p.z, inv, rec.normal.z)};
        if (near_zero(dir)) dir = rec.normal;
        out_o = rec.point;
        out_d = dir;
        att = V3{m.albedo[0], m.albedo[1], m.albedo[2]};
        return true;
    }
    case CRT_MAT_METAL: {        // Material.cuh:86-96 (roughness clamped from above only)
        float rough = m.roughness < 1.f ? m.roughness : 1.f;
        V3 refl = unit_vector(reflect(in_d, rec.normal));
        V3 p = random_in_unit_sphere(rng);
        float inv = 1.0f / sqrtf(length_squared(p));
        V3 u{p.x * inv, p.y * inv, p.z * inv};
        V3 dir{fmaf(rough, u.x, refl.x), fmaf(rough, u.y, refl.y), fmaf(rough, u.z, refl.z)};
        out_o = rec.point;
        out_d = dir;
        att = V3{m.albedo[0], m.albedo[1], m.albedo[2]};
        return dot(dir, rec.normal) > 0;
    }
    case CRT_MAT_DIELECTRIC: {   // Material.cuh:109-137
        att = V3{1.f, 1.f, 1.f};
        float ri = rec.front ? (1.0f / m.ior) : m.ior;
        V3 ud = unit_vector(in_d);
        float cosf32 = fminf(dot_neg(ud, rec.normal), 1.0f);
        double cos_theta = (double)cosf32;
        double sin_theta = sqrt(1.0 - cos_theta * cos_theta);
        bool cannot_refract = (double)ri * sin_theta > 1.0;
        V3 dir;
        bool do_reflect = cannot_refract;
        if (!do_reflect) {
            // schlicksReflectance(float cosine, float ref_idx), Material.cuh:132-137; pow(float,int)
            // resolves to double pow, the sum is one double fma rounded to float.
            float r0 = (1.f - ri) / (1.f + ri);
            r0 = r0 * r0;
            float one_minus = 1.f - cosf32;
            double pw = pow((double)one_minus, 5.0);
            float refl = (float)fma(pw, (double)(1.f - r0), (double)r0);
            float u = xorwow_uniform(rng);   // drawn only when refraction is possible (short-circuit ||)
            do_reflect = refl > u;
        }
        if (do_reflect) dir = reflect(ud, rec.normal);
        else dir = refract(ud, rec.normal, ri);
        out_o = rec.point;
        out_d = dir;
        return true;
    }
    default:                     // DiffuseLight: base-class scatter returns false (Material.cuh:51)
        return false;
    }
}

V3 emit(const crt_material& m) {   // Material.cuh:53-55,144-146
    if (m.type == CRT_MAT_DIFFUSE_LIGHT) return V3{m.emission[0], m.emission[1], m.emission[2]};
    return V3{0.f, 0.f, 0.f};
}

// ---------------------------------------------------------------------------------------------
// camera                                                             reference: Core/Camera.cuh
// ---------------------------------------------------------------------------------------------
struct Cam {
    V3 pos, right, up, ll, horiz, vert;
    float lens_radius;
};

// host-side plain float arithmetic (x86-64 baseline: no FMA contraction), as the reference's
// host compiler evaluates Camera::updateCameraVectors, Camera.cuh:159-182
V3 h_unit(V3 v) {
    float len = sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    float inv = 1 / len;
    return V3{inv * v.x, inv * v.y, inv * v.z};
}
V3 h_cross(V3 u, V3 v) { return V3{u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x}; }

Cam make_camera(const crt_camera& c) {
    const float PI = 3.1415926535897932385f;
    Cam k;
    k.pos = V3{c.position[0], c.position[1], c.position[2]};
    V3 front;
    front.x = -cosf(c.yaw_deg * PI / 180.0f) * cosf(c.pitch_deg * PI / 180.0f);
    front.y = -sinf(c.pitch_deg * PI / 180.0f);
    front.z = -sinf(c.yaw_deg * PI / 180.0f) * cosf(c.pitch_deg * PI / 180.0f);
    front = h_unit(front);
    V3 wup{c.world_up[0], c.world_up[1], c.world_up[2]};
    k.right = h_unit(h_cross(front, wup));
    k.up = h_unit(h_cross(k.right, front));
    float theta = c.vfov_deg * PI / 180.0f;
    float h = tanf(theta / 2.0f);
    float vh = 2.0f * h;
    float vw = c.aspect * vh;
    float sh = c.focus_dist * vw, sv = c.focus_dist * vh;
    k.horiz = V3{sh * k.right.x, sh * k.right.y, sh * k.right.z};
    k.vert = V3{sv * k.up.x, sv * k.up.y, sv * k.up.z};
    // m_Position - m_Horizontal/2 - m_Vertical/2 - m_FocusDist*m_Front ; v/2 = (1/2)*v
    V3 hh{0.5f * k.horiz.x, 0.5f * k.horiz.y, 0.5f * k.horiz.z};
    V3 hv{0.5f * k.vert.x, 0.5f * k.vert.y, 0.5f * k.vert.z};
    V3 ff{c.focus_dist * front.x, c.focus_dist * front.y, c.focus_dist * front.z};
    k.ll = ((k.pos - hh) - hv) - ff;
    k.lens_radius = c.aperture / 2.0f;
    return k;
}

// Camera::getRay, Camera.cuh:32-44 (device; contraction as compiled)
void get_ray(const Cam& k, int px, int py, int W, int H, Xorwow& rng, V3& o, V3& d) {
    float dx, dy;
    while (true) {   // randomPointInUnitDisk, Utility.cuh:55-62
        dx = rand_pm1(rng);
        dy = rand_pm1(rng);
        if (fmaf(dx, dx, dy * dy) >= 1.f) continue;
        break;
    }
    float rdx = k.lens_radius * dx, rdy = k.lens_radius * dy;
    V3 off{fmaf(rdx, k.right.x, rdy * k.up.x), fmaf(rdx, k.right.y, rdy * k.up.y), fmaf(rdx, k.right.z, rdy * k.up.z)};
    float u = (xorwow_uniform(rng) + (float)px) / (float)W;
    float v = (xorwow_uniform(rng) + (float)py) / (float)H;
    o = k.pos + off;
    V3 t{fmaf(v, k.vert.x, fmaf(u, k.horiz.x, k.ll.x)), fmaf(v, k.vert.y, fmaf(u, k.horiz.y, k.ll.y)),
         fmaf(v, k.vert.z, fmaf(u, k.horiz.z, k.ll.z))};
    d = (t - k.pos) - off;
}

// getBackgroundColor, CRTUtility.cuh:34-38
V3 background(V3 d) {
    float inv = 1.0f / sqrtf(length_squared(d));
    float t = fmaf(inv, d.y, 1.0f) * 0.5f;     // 0.5*(unit.y + 1): the scale and +1 fuse
    float omt = 1.0f - t;
    return V3{fmaf(t, 0.5f, omt), fmaf(t, 0.7f, omt), t + omt};
}

// rayColor, CUDAKernels.h:102-145
V3 ray_color(const orc_scene& s, V3 o, V3 d, int max_bounces, Xorwow& rng, int mode, uint64_t& rays) {
    V3 thr{1.f, 1.f, 1.f};
    const int MIN_BOUNCES = 3;
    const float MAX_PROB = 0.95f;
    for (int bounce = 0; bounce < max_bounces; bounce++) {
        if (bounce >= MIN_BOUNCES) {
            float p = fminf(fmaxf(thr.x, fmaxf(thr.y, thr.z)), MAX_PROB);
            if (xorwow_uniform(rng) > p) break;
            float inv = 1.0f / p;                 // operator/= multiplies by 1/t (Vec3.cuh:61-63)
            thr = V3{thr.x * inv, thr.y * inv, thr.z * inv};
        }
        Hit rec;
        rays++;
        if (hit_scene(s, o, d, 0.001f, kInf, rec, mode)) {
            if (rec.mat < (uint32_t)s.materials.size()) {   // uint32 index: ">= 0" is vacuous
                const crt_material& m = s.materials[rec.mat];
                V3 att{0, 0, 0}, no{0, 0, 0}, nd{0, 0, 0};
                if (scatter(m, o, d, rec, att, no, nd, rng)) {
                    thr = thr * att;
                    o = no;
                    d = nd;
                } else {
                    return emit(m);                // raw emission, throughput ignored (:132-134)
                }
            }
            // invalid material index: loop again with the same ray (:127,135)
        } else {
            return thr * background(d);
        }
    }
    return V3{0.f, 0.f, 0.f};
}

// writeColor, CRTUtility.cuh:21-32. linearToGamma(double) on a float: sqrt in double rounded to
// float equals the correctly rounded float sqrt.
uint8_t pack_channel(float c) {
    float g = c > 0 ? (float)sqrt((double)c) : 0.f;
    if (g > 0.999f) g = 0.999f;
    if (g < 0.f) g = 0.f;
    return (uint8_t)(uint32_t)(256 * g);
}

template <class F>
void parallel_rows(int rows, int threads, F fn) {
    if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
    if (threads <= 0) threads = 1;
    threads = std::min(threads, std::max(rows, 1));
    std::atomic<int> next{0};
    std::vector<std::thread> pool;
    auto work = [&]() {
        for (;;) {
            int r = next.fetch_add(1);
            if (r >= rows) break;
            fn(r);
        }
    };
    for (int i = 1; i < threads; i++) pool.emplace_back(work);
    work();
    for (auto& t : pool) t.join();
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// C interface
// ---------------------------------------------------------------------------------------------
extern "C" {

orc_scene* orc_scene_create(const crt_scene* sc) {
    if (!sc) return nullptr;
    if (sc->num_triangles && (!sc->positions || !sc->indices || !sc->tri_material)) return nullptr;
    if (sc->num_spheres && (!sc->spheres || !sc->sphere_material)) return nullptr;
    orc_scene* s = new orc_scene();
    s->num_triangles = (int)sc->num_triangles;
    s->materials.assign(sc->materials, sc->materials + sc->num_materials);
    std::vector<uint32_t> offs;
    if (sc->mesh_tri_offset && sc->num_meshes > 0) offs.assign(sc->mesh_tri_offset, sc->mesh_tri_offset + sc->num_meshes + 1);
    else if (sc->num_triangles > 0) offs = {0u, sc->num_triangles};
    for (size_t mi = 0; mi + 1 < offs.size(); mi++) {
        uint32_t t0 = offs[mi], t1 = offs[mi + 1];
        Mesh m;
        m.first_tri = (int)t0;
        uint32_t vmin = 0xffffffffu, vmax = 0;
        for (uint32_t i = 3 * t0; i < 3 * t1; i++) { vmin = std::min(vmin, sc->indices[i]); vmax = std::max(vmax, sc->indices[i]); }
        if (t1 == t0) { vmin = 0; vmax = 0; }
        for (uint32_t v = vmin; t1 > t0 && v <= vmax; v++)
            m.verts.push_back(V3{sc->positions[3 * v], sc->positions[3 * v + 1], sc->positions[3 * v + 2]});
        for (uint32_t i = 3 * t0; i < 3 * t1; i++) m.indices.push_back(sc->indices[i] - vmin);
        for (uint32_t t = t0; t < t1; t++) { m.face_mat.push_back(sc->tri_material[t]); m.orig_id.push_back((int)t); }
        // Mesh ctor, Mesh.cuh:37-52: bbox over every vertex, then build
        if (!m.verts.empty()) {
            m.bbox = box_empty();
            for (const V3& p : m.verts) box_expand(m.bbox, p);
        }
        build_mesh_bvh(m);
        s->meshes.push_back(std::move(m));
    }
    for (uint32_t j = 0; j < sc->num_spheres; j++) {
        Sphere sp;
        sp.c = V3{sc->spheres[4 * j], sc->spheres[4 * j + 1], sc->spheres[4 * j + 2]};
        sp.r = sc->spheres[4 * j + 3];
        sp.r2 = sp.r * sp.r;
        sp.mat = (uint32_t)sc->sphere_material[j];
        V3 rv{sp.r, sp.r, sp.r};
        sp.bbox = box_from_points(sp.c - rv, sp.c + rv);   // Sphere.cuh:11-16
        s->spheres.push_back(sp);
    }
    for (int i = 0; i < (int)s->meshes.size(); i++) s->objects.push_back(Object{0, i});
    for (int i = 0; i < (int)s->spheres.size(); i++) s->objects.push_back(Object{1, i});
    build_top_bvh(*s);
    return s;
}

void orc_scene_destroy(orc_scene* s) { delete s; }

void orc_scene_permutation(const orc_scene* s, int32_t* perm) {
    for (const Mesh& m : s->meshes)
        for (size_t k = 0; k < m.orig_id.size(); k++) perm[(size_t)m.first_tri + k] = m.orig_id[k];
}

static void load_state(const uint32_t* p, Xorwow& r) { for (int k = 0; k < 5; k++) r.v[k] = p[k]; r.d = p[5]; }
static void store_state(uint32_t* p, const Xorwow& r) { for (int k = 0; k < 5; k++) p[k] = r.v[k]; p[5] = r.d; }

int orc_render(const orc_scene* s, const crt_camera* cam, int W, int H, int x0, int y0, int w, int h,
               int spp, int max_bounces, uint64_t seed, uint64_t subseq_base, uint32_t* rng_state,
               int fresh_rng, int mode, int threads, float* out_radiance, uint8_t* out_rgba8, uint64_t* out_rays) {
    if (!s || !cam || W <= 0 || H <= 0 || w <= 0 || h <= 0 || x0 < 0 || y0 < 0 || x0 + w > W || y0 + h > H || spp <= 0) return -1;
    if (!rng_state && !fresh_rng) return -1;
    Cam k = make_camera(*cam);
    const float scale = 1.f / (float)spp;   // m_PixelSampleScale, Camera.cuh:71
    std::atomic<uint64_t> total_rays{0};
    xw_sequence_table();
    parallel_rows(h, threads, [&](int row) {
        int y = y0 + row;
        uint64_t rays = 0;
        Xorwow first{};
        for (int col = 0; col < w; col++) {
            int x = x0 + col;
            size_t pix = (size_t)y * W + x;
            Xorwow rng;
            if (fresh_rng) {
                // curand_init(seed, subseq_base + pix, 0); consecutive pixels differ by one
                // application of the 2^67 skip matrix
                if (col == 0) first = xorwow_init(seed, subseq_base + pix);
                else xw_apply(xw_sequence_table()[0], first.v);
                rng = first;
            } else {
                load_state(rng_state + pix * 6, rng);
            }
            V3 sum{0.f, 0.f, 0.f};
            for (int sidx = 0; sidx < spp; sidx++) {     // render, CUDAKernels.h:157-162
                V3 o, d;
                get_ray(k, x, y, W, H, rng, o, d);
                V3 c = ray_color(*s, o, d, max_bounces, rng, mode, rays);
                sum = sum + c;
            }
            V3 outc{scale * sum.x, scale * sum.y, scale * sum.z};
            size_t wp = (size_t)row * w + col;
            if (out_radiance) { out_radiance[3 * wp] = outc.x; out_radiance[3 * wp + 1] = outc.y; out_radiance[3 * wp + 2] = outc.z; }
            if (out_rgba8) {
                out_rgba8[4 * wp] = pack_channel(outc.x);
                out_rgba8[4 * wp + 1] = pack_channel(outc.y);
                out_rgba8[4 * wp + 2] = pack_channel(outc.z);
                out_rgba8[4 * wp + 3] = 255;
            }
            if (rng_state) store_state(rng_state + pix * 6, rng);
        }
        total_rays += rays;
    });
    if (out_rays) *out_rays = total_rays.load();
    return 0;
}

int orc_primary(const orc_scene* s, const crt_camera* cam, int W, int H, int x0, int y0, int w, int h,
                uint64_t seed, uint64_t subseq_base, int mode, int threads, int32_t* out_ids, float* out_t) {
    if (!s || !cam || W <= 0 || H <= 0 || w <= 0 || h <= 0 || x0 < 0 || y0 < 0 || x0 + w > W || y0 + h > H) return -1;
    Cam k = make_camera(*cam);
    xw_sequence_table();
    parallel_rows(h, threads, [&](int row) {
        int y = y0 + row;
        Xorwow first{};
        for (int col = 0; col < w; col++) {
            int x = x0 + col;
            size_t pix = (size_t)y * W + x;
            if (col == 0) first = xorwow_init(seed, subseq_base + pix);
            else xw_apply(xw_sequence_table()[0], first.v);
            Xorwow rng = first;
            V3 o, d;
            get_ray(k, x, y, W, H, rng, o, d);
            Hit rec;
            size_t wp = (size_t)row * w + col;
            if (hit_scene(*s, o, d, 0.001f, kInf, rec, mode)) {
                if (out_ids) out_ids[wp] = rec.prim;
                if (out_t) out_t[wp] = rec.t;
            } else {
                if (out_ids) out_ids[wp] = -1;
                if (out_t) out_t[wp] = kInf;
            }
        }
    });
    return 0;
}

void orc_xorwow(uint64_t seed, uint64_t subseq, int n, uint32_t out_state[6], uint32_t* out_raw, float* out_uniform) {
    Xorwow st = xorwow_init(seed, subseq);
    store_state(out_state, st);
    Xorwow a = st, b = st;
    for (int i = 0; i < n; i++) {
        if (out_raw) out_raw[i] = xorwow_next(a);
        if (out_uniform) out_uniform[i] = xorwow_uniform(b);
    }
}

void orc_xorwow_matrix(int k, uint32_t* out800) {
    const XwMatrix& m = xw_sequence_table()[(size_t)k];
    std::memcpy(out800, m.w.data(), 800 * sizeof(uint32_t));
}

void orc_camera_vectors(const crt_camera* cam, float out[19]) {
    Cam k = make_camera(*cam);
    const V3* v[6] = {&k.pos, &k.right, &k.up, &k.ll, &k.horiz, &k.vert};
    for (int i = 0; i < 6; i++) { out[3 * i] = v[i]->x; out[3 * i + 1] = v[i]->y; out[3 * i + 2] = v[i]->z; }
    out[18] = k.lens_radius;
}

void orc_rng_states(uint64_t seed, uint64_t subseq_base, int W, int H, int threads, uint32_t* out6) {
    xw_sequence_table();
    parallel_rows(H, threads, [&](int y) {
        Xorwow first{};
        for (int x = 0; x < W; x++) {
            size_t pix = (size_t)y * W + x;
            if (x == 0) first = xorwow_init(seed, subseq_base + pix);
            else xw_apply(xw_sequence_table()[0], first.v);
            store_state(out6 + pix * 6, first);
        }
    });
}

}  // extern "C"
