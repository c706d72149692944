// crt_bvh_build.h — parallel BVH build: Morton order -> PLOC binary BVH -> 8-wide compressed BVH.
//
// Replaces the reference's two single-GPU-thread builders
// (reference: CudaRayTracer/src/Core/Mesh.cuh:121-219 Mesh::buildBVHMesh,
//  Core/BVHNode.cuh:21-84 BVHNode::buildBVHScene, launched <<<1,1>>> from src/SceneManager.h:95,180).
//
// Every step is a data-parallel functor `void operator()(int i)` run over n items by an execution
// policy. The product library instantiates the builder with the CUDA policy (crt_bvh_build.cu:
// one kernel launch per step, cub radix sort / scan). tests/hostsim instantiates the same functors
// with a sequential host policy to debug the algorithm without a GPU — test infrastructure only.
//
// Algorithm
//   1. primitive boxes + centroid bounds                      (PrimBoxFn)
//   2. 63-bit Morton codes of the centroids, radix sort       (MortonFn, Exec::sort_pairs)
//   3. PLOC (Meister & Bittner 2018): repeat { nearest neighbour within +-R along the Morton
//      order by surface area of the union; merge mutual pairs; compact } until one cluster
//   4. collapse to 8-wide nodes: a bottom-up dynamic programme over the binary tree chooses, for every binary node
//      and every j = 1..7, the cheapest way to cover its subtree with at most j children of a wide node
//      (CollapseCostFn; Ylitie, Karras, Laine 2017, section 3.1), then the wide nodes are emitted level by level
//      following those decisions (CollapseFn): leaves of <= leaf_max primitives, octant-ordered child slots,
//      8-bit quantised child boxes, primitive records written in leaf order.
#pragma once
#include "crt_types.h"

namespace crt {

#if defined(__CUDA_ARCH__)
CRT_HD uint32_t atomic_add_u32(uint32_t* p, uint32_t v) { return atomicAdd(p, v); }
CRT_HD void atomic_min_i32(int32_t* p, int32_t v) { atomicMin(p, v); }
CRT_HD void atomic_max_i32(int32_t* p, int32_t v) { atomicMax(p, v); }
CRT_HD void atomic_max_u64(unsigned long long* p, unsigned long long v) { atomicMax(p, v); }
#else
CRT_HD uint32_t atomic_add_u32(uint32_t* p, uint32_t v) { uint32_t o = *p; *p = o + v; return o; }
CRT_HD void atomic_min_i32(int32_t* p, int32_t v) { if (v < *p) *p = v; }
CRT_HD void atomic_max_i32(int32_t* p, int32_t v) { if (v > *p) *p = v; }
CRT_HD void atomic_max_u64(unsigned long long* p, unsigned long long v) { if (v > *p) *p = v; }
#endif

#if defined(__CUDA_ARCH__)
CRT_HD void thread_fence() { __threadfence(); }
CRT_HD float load_cg(const float* p) { return __ldcg(p); }   // past L1: written by a thread of another SM
#else
CRT_HD void thread_fence() {}
CRT_HD float load_cg(const float* p) { return *p; }
#endif

// order-preserving float <-> int32 (for atomic min/max on floats)
CRT_HD int32_t float_ordered(float f) { int32_t i = (int32_t)f_bits(f); return i >= 0 ? i : i ^ 0x7fffffff; }
CRT_HD float ordered_float(int32_t i) { return bits_f((uint32_t)(i >= 0 ? i : i ^ 0x7fffffff)); }

struct SceneInput {   // device (or host-sim) pointers to the caller's flat arrays
    const float* positions;
    const uint32_t* indices;
    const int32_t* tri_material;
    const float* spheres;
    const int32_t* sphere_material;
    uint32_t num_triangles;
    uint32_t num_spheres;
    const uint32_t* tri_key = nullptr;   // optional: the tie key stored in a triangle's record instead of its id (reference-exact mode)
};

CRT_HD float fmin3(float a, float b, float c) { return fminf(a, fminf(b, c)); }
CRT_HD float fmax3(float a, float b, float c) { return fmaxf(a, fmaxf(b, c)); }

// next representable float towards -inf / +inf (finite inputs)
CRT_HD float f_down(float x) { uint32_t u = f_bits(x); if (x > 0.f) u--; else if (x < 0.f) u++; else u = 0x80000001u; return bits_f(u); }
CRT_HD float f_up(float x) { uint32_t u = f_bits(x); if (x > 0.f) u++; else if (x < 0.f) u--; else u = 0x00000001u; return bits_f(u); }

// ---- step 1: primitive boxes ------------------------------------------------------------
struct PrimBoxFn {
    SceneInput in;
    f4* plo;            // per primitive (lo.xyz, -)
    f4* phi;            // per primitive (hi.xyz, -)
    CRT_HD void operator()(int i) const {
        float lx, ly, lz, hx, hy, hz;
        if ((uint32_t)i < in.num_triangles) {
            const uint32_t a = in.indices[3 * i], b = in.indices[3 * i + 1], c = in.indices[3 * i + 2];
            const float ax = in.positions[3 * a], ay = in.positions[3 * a + 1], az = in.positions[3 * a + 2];
            const float bx = in.positions[3 * b], by = in.positions[3 * b + 1], bz = in.positions[3 * b + 2];
            const float cx = in.positions[3 * c], cy = in.positions[3 * c + 1], cz = in.positions[3 * c + 2];
            lx = fmin3(ax, bx, cx); ly = fmin3(ay, by, cy); lz = fmin3(az, bz, cz);
            hx = fmax3(ax, bx, cx); hy = fmax3(ay, by, cy); hz = fmax3(az, bz, cz);
        } else {
            const uint32_t j = (uint32_t)i - in.num_triangles;
            const float cx = in.spheres[4 * j], cy = in.spheres[4 * j + 1], cz = in.spheres[4 * j + 2];
            const float r = fabsf(in.spheres[4 * j + 3]);
            // c -/+ r rounds to nearest; step one ulp outwards so the box contains the sphere
            lx = f_down(cx - r); ly = f_down(cy - r); lz = f_down(cz - r);
            hx = f_up(cx + r); hy = f_up(cy + r); hz = f_up(cz + r);
        }
        f4 lo; lo.x = lx; lo.y = ly; lo.z = lz; lo.w = 0.f;
        f4 hi; hi.x = hx; hi.y = hy; hi.z = hz; hi.w = 0.f;
        plo[i] = lo;
        phi[i] = hi;
    }
};

// ---- step 1b: giant primitives ------------------------------------------------------------------
// A primitive that dwarfs everything else (the reference always adds a ground sphere of radius 999 under a
// scene of extent 0.6, CUDAKernels.h:69) would be the root's sibling of "all the rest": every ray pays one
// wide-node visit just to step past it, and its centroid stretches the Morton grid. Up to CRT_MAX_GIANTS such
// primitives are taken out of the BVH; every ray tests them directly before it traverses (trace_giants).
#define CRT_MAX_GIANTS 4
#define CRT_GIANT_RATIO 16.0f

struct GiantSet {
    uint32_t count;
    uint32_t id[CRT_MAX_GIANTS];   // ascending
    CRT_HD bool has(uint32_t p) const {
        for (uint32_t k = 0; k < count; k++) if (id[k] == p) return true;
        return false;
    }
    // j-th primitive that is not a giant
    CRT_HD uint32_t remap(uint32_t j) const {
        uint32_t p = j;
        for (uint32_t k = 0; k < count; k++) if (p >= id[k]) p++;
        return p;
    }
};

// largest primitive (by longest box side) outside `skip`: key = extent bits << 32 | prim
struct GiantScanFn {
    const f4* plo;
    const f4* phi;
    GiantSet skip;
    unsigned long long* key;
    CRT_HD void operator()(int i) const {
        if (skip.has((uint32_t)i)) return;
        const float e = fmaxf(phi[i].x - plo[i].x, fmaxf(phi[i].y - plo[i].y, phi[i].z - plo[i].z));
        atomic_max_u64(key, ((unsigned long long)f_bits(e > 0.f ? e : 0.f) << 32) | (unsigned long long)(uint32_t)i);
    }
};

// bounds of all primitive boxes outside `skip` (float_ordered min xyz, max xyz)
struct RestBoundsFn {
    const f4* plo;
    const f4* phi;
    GiantSet skip;
    int32_t* bounds;
    CRT_HD void operator()(int i) const {
        if (skip.has((uint32_t)i)) return;
        atomic_min_i32(&bounds[0], float_ordered(plo[i].x));
        atomic_min_i32(&bounds[1], float_ordered(plo[i].y));
        atomic_min_i32(&bounds[2], float_ordered(plo[i].z));
        atomic_max_i32(&bounds[3], float_ordered(phi[i].x));
        atomic_max_i32(&bounds[4], float_ordered(phi[i].y));
        atomic_max_i32(&bounds[5], float_ordered(phi[i].z));
    }
};

// centroid bounds of the primitives that stay in the BVH
struct CentroidBoundsFn {
    const f4* plo;
    const f4* phi;
    GiantSet giants;
    int32_t* cbounds;   // centroid bounds, float_ordered: min xyz, max xyz
    CRT_HD void operator()(int j) const {
        const uint32_t p = giants.remap((uint32_t)j);
        const float mx = 0.5f * (plo[p].x + phi[p].x), my = 0.5f * (plo[p].y + phi[p].y), mz = 0.5f * (plo[p].z + phi[p].z);
        atomic_min_i32(&cbounds[0], float_ordered(mx));
        atomic_min_i32(&cbounds[1], float_ordered(my));
        atomic_min_i32(&cbounds[2], float_ordered(mz));
        atomic_max_i32(&cbounds[3], float_ordered(mx));
        atomic_max_i32(&cbounds[4], float_ordered(my));
        atomic_max_i32(&cbounds[5], float_ordered(mz));
    }
};

struct GiantRecordsFn {   // primitive records of the giants, appended after the BVH's records
    SceneInput in;
    GiantSet giants;
    PrimRec* prims;
    uint32_t base;
    CRT_HD void operator()(int k) const;
};

// ---- step 2: Morton codes -----------------------------------------------------------------
CRT_HD uint64_t spread21(uint32_t v) {   // 21 bits -> every third bit of 63
    uint64_t x = v & 0x1fffffu;
    x = (x | (x << 32)) & 0x1f00000000ffffull;
    x = (x | (x << 16)) & 0x1f0000ff0000ffull;
    x = (x | (x << 8)) & 0x100f00f00f00f00full;
    x = (x | (x << 4)) & 0x10c30c30c30c30c3ull;
    x = (x | (x << 2)) & 0x1249249249249249ull;
    return x;
}

struct MortonFn {
    const f4* plo;
    const f4* phi;
    const int32_t* cbounds;
    GiantSet giants;
    uint64_t* keys;
    uint32_t* vals;     // primitive number (scene numbering)
    CRT_HD void operator()(int j) const {
        const uint32_t i = giants.remap((uint32_t)j);
        const float bx = ordered_float(cbounds[0]), by = ordered_float(cbounds[1]), bz = ordered_float(cbounds[2]);
        const float ex = ordered_float(cbounds[3]) - bx, ey = ordered_float(cbounds[4]) - by, ez = ordered_float(cbounds[5]) - bz;
        const float mx = 0.5f * (plo[i].x + phi[i].x), my = 0.5f * (plo[i].y + phi[i].y), mz = 0.5f * (plo[i].z + phi[i].z);
        const float scale = 2097151.0f;
        const float nx = ex > 0.f ? (mx - bx) / ex : 0.f, ny = ey > 0.f ? (my - by) / ey : 0.f, nz = ez > 0.f ? (mz - bz) / ez : 0.f;
        const uint32_t qx = (uint32_t)fminf(fmaxf(nx * scale, 0.f), scale);
        const uint32_t qy = (uint32_t)fminf(fmaxf(ny * scale, 0.f), scale);
        const uint32_t qz = (uint32_t)fminf(fmaxf(nz * scale, 0.f), scale);
        keys[j] = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
        vals[j] = i;
    }
};

// ---- binary BVH storage: node k box = (blo[k].xyz, bhi[k].xyz); blo[k].w = left child,
// bhi[k].w = right child (as uint bits). Leaf: left = 0xffffffff, right = primitive number.
// Nodes [0,N) are the leaves in Morton order, inner nodes are appended from N.
CRT_HD f4 mk_f4(float x, float y, float z, uint32_t wbits) { f4 r; r.x = x; r.y = y; r.z = z; r.w = bits_f(wbits); return r; }

struct InitLeavesFn {
    const f4* plo;
    const f4* phi;
    const uint32_t* sorted_prim;
    f4* blo;
    f4* bhi;
    uint32_t* bcount;
    uint32_t* cluster;
    CRT_HD void operator()(int i) const {
        const uint32_t p = sorted_prim[i];
        blo[i] = mk_f4(plo[p].x, plo[p].y, plo[p].z, 0xffffffffu);
        bhi[i] = mk_f4(phi[p].x, phi[p].y, phi[p].z, p);
        bcount[i] = 1u;
        cluster[i] = (uint32_t)i;
    }
};


// ---- step 3: PLOC ---------------------------------------------------------------------------
#ifndef CRT_PLOC_RADIUS
#define CRT_PLOC_RADIUS 24
#endif

CRT_HD float union_area(const f4& alo, const f4& ahi, const f4& blo_, const f4& bhi_) {
    const float ex = fmaxf(ahi.x, bhi_.x) - fminf(alo.x, blo_.x);
    const float ey = fmaxf(ahi.y, bhi_.y) - fminf(alo.y, blo_.y);
    const float ez = fmaxf(ahi.z, bhi_.z) - fminf(alo.z, blo_.z);
    return ex * ey + ey * ez + ez * ex;
}

// nearest neighbour of cluster i among clusters [i-R, i+R]; ties -> smallest index (this makes
// the lexicographically smallest closest pair mutual, so every round merges at least one pair)
struct PlocNearestFn {
    const f4* blo;
    const f4* bhi;
    const uint32_t* cluster;
    uint32_t* nearest;
    int n;
    CRT_HD void operator()(int i) const {
        const uint32_t ci = cluster[i];
        const f4 lo = blo[ci], hi = bhi[ci];
        const int j0 = i - CRT_PLOC_RADIUS < 0 ? 0 : i - CRT_PLOC_RADIUS;
        const int j1 = i + CRT_PLOC_RADIUS > n - 1 ? n - 1 : i + CRT_PLOC_RADIUS;
        float best = INFINITY;
        int bj = -1;
        for (int j = j0; j <= j1; j++) {
            if (j == i) continue;
            const uint32_t cj = cluster[j];
            const float a = union_area(lo, hi, blo[cj], bhi[cj]);
            if (a < best || bj < 0) { best = a; bj = j; }
        }
        nearest[i] = (uint32_t)bj;
    }
};

// merge mutual nearest pairs; merged[i] = new cluster id, 0xffffffff when i was absorbed
struct PlocMergeFn {
    f4* blo;
    f4* bhi;
    uint32_t* bcount;
    const uint32_t* cluster;
    const uint32_t* nearest;
    uint32_t* merged;
    uint32_t* flags;
    uint32_t* node_counter;   // number of binary nodes allocated so far
    CRT_HD void operator()(int i) const {
        const uint32_t j = nearest[i];
        uint32_t out = cluster[i];
        if (nearest[j] == (uint32_t)i) {
            if ((uint32_t)i < j) {
                const uint32_t a = cluster[i], b = cluster[j];
                const uint32_t id = atomic_add_u32(node_counter, 1u);
                const f4 alo = blo[a], ahi = bhi[a], blo_ = blo[b], bhi_ = bhi[b];
                blo[id] = mk_f4(fminf(alo.x, blo_.x), fminf(alo.y, blo_.y), fminf(alo.z, blo_.z), a);
                bhi[id] = mk_f4(fmaxf(ahi.x, bhi_.x), fmaxf(ahi.y, bhi_.y), fmaxf(ahi.z, bhi_.z), b);
                bcount[id] = bcount[a] + bcount[b];
                out = id;
            } else {
                out = 0xffffffffu;
            }
        }
        merged[i] = out;
        flags[i] = out != 0xffffffffu ? 1u : 0u;
    }
};

// scatter survivors to their scanned positions (keeps Morton order); item n-1 publishes the total
struct PlocCompactFn {
    const uint32_t* merged;
    const uint32_t* flags;
    const uint32_t* pos;      // exclusive scan of flags
    uint32_t* cluster_out;
    uint32_t* total;
    int n;
    CRT_HD void operator()(int i) const {
        if (flags[i]) cluster_out[pos[i]] = merged[i];
        if (i == n - 1) *total = pos[i] + flags[i];
    }
};


// ---- step 4: collapse to the 8-wide compressed layout ---------------------------------------
CRT_HD float box_max_extent(const f4& lo, const f4& hi) { return fmaxf(hi.x - lo.x, fmaxf(hi.y - lo.y, hi.z - lo.z)); }
CRT_HD float box_half_area(const f4& lo, const f4& hi) {
    const float ex = hi.x - lo.x, ey = hi.y - lo.y, ez = hi.z - lo.z;
    return ex * ey + ey * ez + ez * ex;
}

// exponent byte e such that 2^(e-127) * 255 >= extent  (0 for a zero extent: scale 0)
CRT_HD uint32_t quant_exponent(float extent) {
    if (!(extent > 0.f)) return 0u;
    int e;
    const double m = frexp((double)extent / 255.0, &e);   // extent/255 = m * 2^e, m in [0.5,1)
    if (m == 0.5) e -= 1;                                  // exact power of two
    int biased = e + 127;
    if (biased < 1) biased = 1;
    if (biased > 254) biased = 254;
    return (uint32_t)biased;
}

CRT_HD PrimRec make_prim_record(const SceneInput& in, uint32_t p) {
    PrimRec r;
    if (p < in.num_triangles) {
        const uint32_t a = in.indices[3 * p], b = in.indices[3 * p + 1], c = in.indices[3 * p + 2];
        const vec3 v0 = mk3(in.positions[3 * a], in.positions[3 * a + 1], in.positions[3 * a + 2]);
        const vec3 v1 = mk3(in.positions[3 * b], in.positions[3 * b + 1], in.positions[3 * b + 2]);
        const vec3 v2 = mk3(in.positions[3 * c], in.positions[3 * c + 1], in.positions[3 * c + 2]);
        const vec3 e1 = v_sub(v1, v0), e2 = v_sub(v2, v0);   // Mesh.cuh:277-278
        r.q0.x = v0.x; r.q0.y = v0.y; r.q0.z = v0.z; r.q0.w = e1.x;
        r.q1.x = e1.y; r.q1.y = e1.z; r.q1.z = e2.x; r.q1.w = e2.y;
        r.q2.x = e2.z; r.q2.y = bits_f(in.tri_key ? in.tri_key[p] : p); r.q2.z = bits_f((uint32_t)in.tri_material[p]); r.q2.w = bits_f((uint32_t)PRIM_TRIANGLE);
    } else {
        const uint32_t j = p - in.num_triangles;
        const float rad = in.spheres[4 * j + 3];
        r.q0.x = in.spheres[4 * j]; r.q0.y = in.spheres[4 * j + 1]; r.q0.z = in.spheres[4 * j + 2]; r.q0.w = rad;
        r.q1.x = f_mul(rad, rad); r.q1.y = 0.f; r.q1.z = 0.f; r.q1.w = 0.f;   // m_RadSquared, Sphere.cuh:12
        r.q2.x = 0.f; r.q2.y = bits_f(p); r.q2.z = bits_f((uint32_t)in.sphere_material[j]); r.q2.w = bits_f((uint32_t)PRIM_SPHERE);
    }
    return r;
}

CRT_HD void GiantRecordsFn::operator()(int k) const { prims[base + (uint32_t)k] = make_prim_record(in, giants.id[k]); }

// ---- collapse, step 1: which binary subtrees become the children of which wide node ----------------------------
// cost(n, 1) = min( leaf: area * prims * C_PRIM   (only if the subtree holds <= leaf_max primitives),
//                   wide node rooted at n: min_i cost(left, i) + cost(right, 8 - i) + area * C_NODE )
// cost(n, j) = min( cost(n, j - 1), min_k cost(left, k) + cost(right, j - k) )        j = 2..7
// (area = half surface area of the subtree's box: the SAH's hit probability up to a constant). The greedy "open the
// largest child until there are 8" rule that this replaces left the wide nodes half empty (3.9 of 8 slots used on
// the benchmark scenes: subtrees of 3-5 primitives at the bottom of a PLOC tree each became a node of their own).
// C_PRIM / C_NODE = 0.7: measured issue cost of a primitive test (60 instructions at ~9 active lanes) against a node
// visit (250 instructions at ~27 lanes) in k_wf_extend.
#ifndef CRT_COST_NODE
#define CRT_COST_NODE 1.0f
#endif
#ifndef CRT_COST_PRIM
#define CRT_COST_PRIM 0.7f
#endif

struct ParentFn {   // over the inner binary nodes [first_inner, first_inner + n)
    const f4* blo;
    const f4* bhi;
    uint32_t* parent;
    uint32_t first_inner;
    CRT_HD void operator()(int k) const {
        const uint32_t b = first_inner + (uint32_t)k;
        parent[f_bits(blo[b].w)] = b;
        parent[f_bits(bhi[b].w)] = b;
    }
};

struct CollapseCostFn {   // over the leaves; a thread climbs while it is the second child to arrive at a node
    const f4* blo;
    const f4* bhi;
    const uint32_t* bcount;
    const uint32_t* parent;
    uint32_t* arrived;     // zero-filled
    float* cost;           // [node * 8 + j], j = 1..7
    uint8_t* dec;          // [node * 8 + j]: j = 1: 0 = leaf, i = wide node with i children from the left subtree;
                           //                  j > 1: 0 = as cost(n, j - 1), k = k children from the left, j - k from the right
    uint32_t leaf_max;
    uint32_t root;

    CRT_HD void operator()(int i) const {
        uint32_t n = (uint32_t)i;
        {
            const float a = box_half_area(blo[n], bhi[n]) * CRT_COST_PRIM;
            for (int j = 1; j < 8; j++) { cost[(size_t)n * 8 + j] = a; dec[(size_t)n * 8 + j] = 0; }
        }
        while (n != root) {
            const uint32_t p = parent[n];
            thread_fence();
            if (atomic_add_u32(&arrived[p], 1u) == 0u) return;   // the sibling's thread goes on from here
            thread_fence();
            const uint32_t l = f_bits(blo[p].w), r = f_bits(bhi[p].w);
            float cl[8], cr[8];
            for (int j = 1; j < 8; j++) { cl[j] = load_cg(&cost[(size_t)l * 8 + j]); cr[j] = load_cg(&cost[(size_t)r * 8 + j]); }
            const float area = box_half_area(blo[p], bhi[p]);
            float best = INFINITY;
            int best_i = 1;
            for (int k = 1; k < 8; k++) {
                const float v = cl[k] + cr[8 - k];
                if (v < best) { best = v; best_i = k; }
            }
            best += area * CRT_COST_NODE;
            float c[8];
            c[1] = best;
            uint8_t d1 = (uint8_t)best_i;
            if (bcount[p] <= leaf_max) {
                const float leaf = area * (float)bcount[p] * CRT_COST_PRIM;
                if (leaf <= best) { c[1] = leaf; d1 = 0; }
            }
            dec[(size_t)p * 8 + 1] = d1;
            for (int j = 2; j < 8; j++) {
                float b = c[j - 1];
                uint8_t d = 0;
                for (int k = 1; k < j; k++) {
                    const float v = cl[k] + cr[j - k];
                    if (v < b) { b = v; d = (uint8_t)k; }
                }
                c[j] = b;
                dec[(size_t)p * 8 + j] = d;
            }
            for (int j = 1; j < 8; j++) cost[(size_t)p * 8 + j] = c[j];
            n = p;
        }
    }
};

struct CollapseFn {
    const f4* blo;
    const f4* bhi;
    const uint32_t* bcount;
    SceneInput in;
    WideNode* nodes;
    PrimRec* prims;
    uint32_t* wide_bnode;    // binary node collapsed into wide node k
    uint32_t* counters;      // [0] wide nodes allocated, [1] primitive records allocated
    uint32_t level_begin;
    uint32_t leaf_max;       // <= 3
    const uint8_t* dec;      // CollapseCostFn's decisions

    CRT_HD uint32_t left_of(uint32_t b) const { return f_bits(blo[b].w); }
    CRT_HD uint32_t right_of(uint32_t b) const { return f_bits(bhi[b].w); }

    CRT_HD void operator()(int k) const {
        const uint32_t w = level_begin + (uint32_t)k;
        const uint32_t b = wide_bnode[w];
        uint32_t ch[8];
        int nch;
        // node frame
        const f4 nlo = blo[b], nhi = bhi[b];
        // Child boxes are stored on this node's 8-bit grid (step = extent/255 per axis). A subtree that is
        // tiny next to the frame (the Cornell box beside the reference's r=999 ground sphere,
        // CUDAKernels.h:69) would be inflated to a whole grid cell and hit by every ray, so such a
        // subtree is neither opened into this node nor stored as a leaf slot: it stays an inner child
        // and gets a wide node with its own tight frame.
        const float frame = fmax3(nhi.x - nlo.x, nhi.y - nlo.y, nhi.z - nlo.z);
        const float open_min = frame * (1.0f / 16.0f);     // open a child only if it spans >= 16 grid steps
        const float leaf_min = frame * (1.0f / 64.0f);     // leaf slot only if it spans >= 4 grid steps
        if (bcount[b] <= leaf_max) {          // a subtree of <= leaf_max primitives: one leaf slot
            ch[0] = b;
            nch = 1;
        } else {
            // follow the dynamic programme: this node takes d children from its left subtree and 8 - d from its right;
            // a subtree asked for j > 1 children is split again as the programme decided for (subtree, j)
            uint32_t pend_n[8];
            int pend_j[8], np = 0;
            nch = 0;
            {
                const int d = (int)dec[(size_t)b * 8 + 1];
                const int dl = d >= 1 && d <= 7 ? d : 4;
                pend_n[np] = right_of(b); pend_j[np++] = 8 - dl;
                pend_n[np] = left_of(b); pend_j[np++] = dl;
            }
            while (np > 0) {
                const uint32_t n = pend_n[--np];
                int j = pend_j[np];
                int d = 0;
                // binary leaves, subtrees the programme keeps whole, and subtrees that are tiny next to this frame
                // (they get a frame of their own, see above) are children as they are
                if (left_of(n) != 0xffffffffu && box_max_extent(blo[n], bhi[n]) >= open_min) {
                    while (j > 1 && (d = (int)dec[(size_t)n * 8 + j]) == 0) j--;
                } else {
                    j = 1;
                }
                if (j <= 1) { ch[nch++] = n; continue; }
                pend_n[np] = right_of(n); pend_j[np++] = j - d;
                pend_n[np] = left_of(n); pend_j[np++] = d;
            }
        }
        // leaf slot or inner child?
        bool as_leaf[8];
        for (int i = 0; i < nch; i++)
            as_leaf[i] = bcount[ch[i]] <= leaf_max && (nch == 1 || box_max_extent(blo[ch[i]], bhi[ch[i]]) >= leaf_min);
        const uint32_t ex = quant_exponent(nhi.x - nlo.x), ey = quant_exponent(nhi.y - nlo.y), ez = quant_exponent(nhi.z - nlo.z);
        const double scx = ex ? ldexp(1.0, (int)ex - 127) : 0.0, scy = ey ? ldexp(1.0, (int)ey - 127) : 0.0, scz = ez ? ldexp(1.0, (int)ez - 127) : 0.0;
        const float ncx = 0.5f * (nlo.x + nhi.x), ncy = 0.5f * (nlo.y + nhi.y), ncz = 0.5f * (nlo.z + nhi.z);

        // greedy child -> slot assignment: slot s is visited first by rays whose direction sign
        // per axis is + where bit (4,2,1) of s is 0, so it should hold the child lying furthest
        // towards - on those axes: cost(c,s) = dot(centroid_c - centroid_node, sign(s))
        int slot_child[8];
        for (int s = 0; s < 8; s++) slot_child[s] = -1;
        bool used[8];
        for (int i = 0; i < 8; i++) used[i] = false;
        float cost[8][8];
        for (int i = 0; i < nch; i++) {
            const f4 lo = blo[ch[i]], hi = bhi[ch[i]];
            const float dx = 0.5f * (lo.x + hi.x) - ncx, dy = 0.5f * (lo.y + hi.y) - ncy, dz = 0.5f * (lo.z + hi.z) - ncz;
            for (int s = 0; s < 8; s++) {
                const float sxn = (s & 4) ? 1.f : -1.f, syn = (s & 2) ? 1.f : -1.f, szn = (s & 1) ? 1.f : -1.f;
                cost[i][s] = dx * sxn + dy * syn + dz * szn;
            }
        }
        for (int it = 0; it < nch; it++) {
            float bc = -INFINITY;
            int bi = -1, bs = -1;
            for (int i = 0; i < nch; i++) {
                if (used[i]) continue;
                for (int s = 0; s < 8; s++) {
                    if (slot_child[s] >= 0) continue;
                    if (bi < 0 || cost[i][s] > bc) { bc = cost[i][s]; bi = i; bs = s; }
                }
            }
            used[bi] = true;
            slot_child[bs] = bi;
        }
        // allocate children
        uint32_t n_inner = 0, n_prims = 0;
        for (int i = 0; i < nch; i++) {
            if (as_leaf[i]) n_prims += bcount[ch[i]];
            else n_inner++;
        }
        const uint32_t child_base = n_inner ? atomic_add_u32(&counters[0], n_inner) : 0u;
        const uint32_t prim_base = n_prims ? atomic_add_u32(&counters[1], n_prims) : 0u;

        uint32_t imask = 0u;
        uint32_t meta[8], qlx[8], qly[8], qlz[8], qhx[8], qhy[8], qhz[8];
        uint32_t inner_rank = 0, prim_off = 0;
        for (int s = 0; s < 8; s++) {
            meta[s] = 0u;
            qlx[s] = qly[s] = qlz[s] = 255u;      // empty slot: inverted box, never hit
            qhx[s] = qhy[s] = qhz[s] = 0u;
            const int i = slot_child[s];
            if (i < 0) continue;
            const uint32_t c = ch[i];
            const f4 lo = blo[c], hi = bhi[c];
            // exact in double: floor/ceil of (bound - p) / 2^e
            double v;
            v = scx > 0.0 ? floor(((double)lo.x - (double)nlo.x) / scx) : 0.0; qlx[s] = (uint32_t)(v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v));
            v = scy > 0.0 ? floor(((double)lo.y - (double)nlo.y) / scy) : 0.0; qly[s] = (uint32_t)(v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v));
            v = scz > 0.0 ? floor(((double)lo.z - (double)nlo.z) / scz) : 0.0; qlz[s] = (uint32_t)(v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v));
            v = scx > 0.0 ? ceil(((double)hi.x - (double)nlo.x) / scx) : 0.0; qhx[s] = (uint32_t)(v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v));
            v = scy > 0.0 ? ceil(((double)hi.y - (double)nlo.y) / scy) : 0.0; qhy[s] = (uint32_t)(v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v));
            v = scz > 0.0 ? ceil(((double)hi.z - (double)nlo.z) / scz) : 0.0; qhz[s] = (uint32_t)(v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v));
            if (!as_leaf[i]) {
                imask |= 1u << s;
                meta[s] = (1u << 5) | (24u + (uint32_t)s);
                wide_bnode[child_base + inner_rank] = c;
                inner_rank++;
            } else {
                // gather the <= 3 primitives of this subtree (depth-first, left first)
                const uint32_t cnt = bcount[c];
                uint32_t st[4];
                int sp = 0;
                st[sp++] = c;
                uint32_t written = 0;
                while (sp > 0) {
                    const uint32_t x = st[--sp];
                    if (left_of(x) == 0xffffffffu) {
                        prims[prim_base + prim_off + written] = make_prim_record(in, right_of(x));
                        written++;
                    } else {
                        st[sp++] = right_of(x);
                        st[sp++] = left_of(x);
                    }
                }
                const uint32_t unary = cnt >= 3 ? 7u : (cnt == 2 ? 3u : 1u);
                meta[s] = (unary << 5) | prim_off;
                prim_off += cnt;
            }
        }
        WideNode out;
        out.n0.x = f_bits(nlo.x); out.n0.y = f_bits(nlo.y); out.n0.z = f_bits(nlo.z);
        out.n0.w = ex | (ey << 8) | (ez << 16) | (imask << 24);
        out.n1.x = child_base; out.n1.y = prim_base;
        out.n1.z = meta[0] | (meta[1] << 8) | (meta[2] << 16) | (meta[3] << 24);
        out.n1.w = meta[4] | (meta[5] << 8) | (meta[6] << 16) | (meta[7] << 24);
#define CRT_PACK4(a, o) ((a)[o] | ((a)[o + 1] << 8) | ((a)[o + 2] << 16) | ((a)[o + 3] << 24))
        out.n2.x = CRT_PACK4(qlx, 0); out.n2.y = CRT_PACK4(qlx, 4); out.n2.z = CRT_PACK4(qly, 0); out.n2.w = CRT_PACK4(qly, 4);
        out.n3.x = CRT_PACK4(qlz, 0); out.n3.y = CRT_PACK4(qlz, 4); out.n3.z = CRT_PACK4(qhx, 0); out.n3.w = CRT_PACK4(qhx, 4);
        out.n4.x = CRT_PACK4(qhy, 0); out.n4.y = CRT_PACK4(qhy, 4); out.n4.z = CRT_PACK4(qhz, 0); out.n4.w = CRT_PACK4(qhz, 4);
#undef CRT_PACK4
        nodes[w] = out;
    }
};

}  // namespace crt
