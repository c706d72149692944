// crt_trace.h — closest-hit query: compressed 8-wide BVH traversal + primitive tests.
//
// Replaces BVHNode::hit -> virtual Mesh::hit / Sphere::hit -> AABB::hit
// (reference: CudaRayTracer/src/Core/BVHNode.cuh:115-156, Core/Mesh.cuh:55-110,266-308,
// Core/Sphere.cuh:18-47, Core/AABB.cuh:123-146).
//
// What must match the reference bit-for-bit is the set of accepted primitives and their t:
//   - triangle: Moeller-Trumbore with eps 1e-8, inclusive edges, t in [tmin, closest]   (Mesh.cuh:266-293)
//   - sphere  : half-b quadratic, nearer root in range else farther                     (Sphere.cuh:18-34)
// both evaluated with the reference's compiled rounding sequence (crt_math.h).
// Box tests only have to be conservative: a different BVH may visit primitives in a different
// order, so the reference's "later equal-t hit replaces" (Mesh.cuh:293) is made order-independent
// here: among equal t the larger primitive id wins.
#pragma once
#include "crt_types.h"

namespace crt {

struct HitRec {
    float t;
    uint32_t prim;   // index into SceneView::prims (leaf order)
    int32_t id;      // caller's primitive id, -1 = miss
};

struct TraceCounters {
    uint32_t nodes;
    uint32_t prims;
};

#ifndef CRT_STACK_ENTRIES
#define CRT_STACK_ENTRIES 48
#endif

// L1 eviction priority of the scene loads (experiment): CRT_NODE_L1_HINT 1 = wide nodes evict_last;
// CRT_PRIM_L1_HINT 1 = primitive records evict_first, 2 = no_allocate. 0 = plain ld.global.nc (__ldg).
#ifndef CRT_NODE_L1_HINT
#define CRT_NODE_L1_HINT 0
#endif
#ifndef CRT_PRIM_L1_HINT
#define CRT_PRIM_L1_HINT 0
#endif
#if defined(__CUDA_ARCH__)
CRT_HD int hi_bit(uint32_t v) { return 31 - __clz((int)v); }
CRT_HD int pop_count(uint32_t v) { return __popc(v); }
CRT_HD f4 ld_f4(const f4* p) {
    f4 r;
#if CRT_PRIM_L1_HINT == 1
    asm("ld.global.nc.L1::evict_first.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
#elif CRT_PRIM_L1_HINT == 2
    asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
#else
    float4 v = __ldg(reinterpret_cast<const float4*>(p)); r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w;
#endif
    return r;
}
CRT_HD u4 ld_u4(const u4* p) {
    u4 r;
#if CRT_NODE_L1_HINT == 1
    asm("ld.global.nc.L1::evict_last.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
#else
    uint4 v = __ldg(reinterpret_cast<const uint4*>(p)); r.x = v.x; r.y = v.y; r.z = v.z; r.w = v.w;
#endif
    return r;
}
#else
CRT_HD int hi_bit(uint32_t v) { return 31 - __builtin_clz(v); }
CRT_HD int pop_count(uint32_t v) { return __builtin_popcount(v); }
CRT_HD f4 ld_f4(const f4* p) { return *p; }
CRT_HD u4 ld_u4(const u4* p) { return *p; }
#endif

CRT_HD float byte_f(uint32_t w, int i) { return (float)((w >> (8 * i)) & 0xffu); }
// 32768 + byte i of w, exactly, built with ONE byte permute instead of shift + mask + int-to-float:
// 0x47000000 is 32768.0f and mantissa bit 8 weighs 1.0 at that exponent, so dropping the byte into bits 8..15
// adds it. The child-box bounds are then t = f*s + (o - 32768*s); the extra rounding of (o - 32768*s) is at
// most 2^-9 |s| (1/512 of a grid step) and is covered by the conservative pad in visit_node.
#define CRT_BYTE_MAGIC 32768.0f
#if defined(__CUDACC__)
static __device__ uint32_t g_byte_magic = 0x47000000u;   // see load_magic
#endif
#if defined(__CUDA_ARCH__)
// `magic` is 0x47000000 in a register whose value ptxas cannot know (load_magic reads it from a __device__ word):
// given the literal, ptxas puts it into PRMT's one immediate slot and then spends a MOV per permute on bringing the
// selector into a register (43 of the 296 instructions of a node visit in the SASS of the first version).
CRT_HD float byte_magic(uint32_t w, int i, uint32_t magic) { return __uint_as_float(__byte_perm(w, magic, 0x7504u | ((uint32_t)i << 4))); }
CRT_HD uint32_t load_magic() { return *(const volatile uint32_t*)&g_byte_magic; }
#else
CRT_HD float byte_magic(uint32_t w, int i, uint32_t) { return bits_f(0x47000000u | (((w >> (8 * i)) & 0xffu) << 8)); }
CRT_HD uint32_t load_magic() { return 0x47000000u; }
#endif

// Primitive test against the interval [tmin, tmax] (both inclusive, Mesh.cuh:293 / Interval.cuh:33-35).
// Returns the hit parameter and the caller's primitive id; no tie rule here.
struct PrimLoad {   // the three 16-byte loads of a primitive record, issued before the arithmetic that needs them
    f4 q0, q1, q2;
};
CRT_HD PrimLoad load_prim(const PrimRec* prims, uint32_t index) {
    PrimLoad r;
    r.q0 = ld_f4(&prims[index].q0);
    r.q1 = ld_f4(&prims[index].q1);
    r.q2 = ld_f4(&prims[index].q2);
    return r;
}
CRT_HD bool prim_hit_loaded(const PrimLoad& pr, vec3 o, vec3 d, float tmin, float tmax, float& t_out, int32_t& id_out) {
    const f4 q0 = pr.q0, q1 = pr.q1, q2 = pr.q2;
    float t;
    if (f_bits(q2.w) == (uint32_t)PRIM_TRIANGLE) {
        // Mesh::rayTriangleIntersect, Mesh.cuh:266-293
        const vec3 v0 = mk3(q0.x, q0.y, q0.z);
        const vec3 e1 = mk3(q0.w, q1.x, q1.y);
        const vec3 e2 = mk3(q1.z, q1.w, q2.x);
        const vec3 h = v_cross(d, e2);
        const float a = v_dot(e1, h);
        if (fabsf(a) < 1e-8f) return false;
        const float f = f_rcp(a);
        const vec3 s = v_sub(o, v0);
        const float u = f_mul(f, v_dot(s, h));
        if (u < 0.f || u > 1.f) return false;
        const vec3 q = v_cross(s, e1);
        const float v = f_mul(f, v_dot(d, q));
        if (v < 0.f || f_add(u, v) > 1.f) return false;
        t = f_mul(f, v_dot(e2, q));
        if (t < tmin || t > tmax) return false;
    } else {
        // Sphere::hit, Sphere.cuh:18-34
        const vec3 oc = v_sub(o, mk3(q0.x, q0.y, q0.z));
        const float a = v_dot(d, d);
        const float half_b = v_dot(oc, d);
        const float c = f_sub(v_dot(oc, oc), q1.x);
        const float disc = f_fma(half_b, half_b, -f_mul(a, c));
        if (disc < 0.f) return false;
        const float sq = f_sqrt(disc);
        t = f_div(f_sub(-half_b, sq), a);
        if (t < tmin || t > tmax) {
            t = f_div(f_sub(sq, half_b), a);
            if (t < tmin || t > tmax) return false;
        }
    }
    t_out = t;
    id_out = (int32_t)f_bits(q2.y);
    return true;
}

CRT_HD bool prim_hit(const PrimRec* prims, uint32_t index, vec3 o, vec3 d, float tmin, float tmax, float& t_out, int32_t& id_out) {
    return prim_hit_loaded(load_prim(prims, index), o, d, tmin, tmax, t_out, id_out);
}

// Acceptance rule on an already loaded record (see test_prim).
CRT_HD bool test_prim_loaded(const PrimLoad& pr, uint32_t index, vec3 o, vec3 d, float tmin, HitRec& rec) {
    float t;
    int32_t id;
    if (!prim_hit_loaded(pr, o, d, tmin, rec.t, t, id)) return false;
    if (t == rec.t && id < rec.id) return false;
    rec.t = t;
    rec.prim = index;
    rec.id = id;
    return true;
}

// Sequential form: returns true when `rec` accepts the hit and updates rec.
CRT_HD bool test_prim(const PrimRec* prims, uint32_t index, vec3 o, vec3 d, float tmin, HitRec& rec) {
    float t;
    int32_t id;
    if (!prim_hit(prims, index, o, d, tmin, rec.t, t, id)) return false;
    if (t == rec.t && id < rec.id) return false;   // order-independent tie rule
    rec.t = t;
    rec.prim = index;
    rec.id = id;
    return true;
}

// One wide-node visit: the 8 quantised child boxes against the ray interval [tmin, tmax].
// Out: ng = node group (first child index, hit bits [31:24] | inner mask [7:0]),
//      tg = primitive group (first primitive index, hit bits [23:0]). Hit bits are ordered so that the
//      highest bit is the child the ray enters first (octant order).
CRT_HD void visit_node(const WideNode* node, vec3 o, float idx, float idy, float idz, float tmin, float tmax, uint32_t octinv,
                       uint32_t& ng_x, uint32_t& ng_y, uint32_t& tg_x, uint32_t& tg_y, uint32_t mg) {
    const bool negx = idx < 0.f, negy = idy < 0.f, negz = idz < 0.f;
    const uint32_t octinv4 = octinv * 0x01010101u;
    const u4 n0 = ld_u4(&node->n0);
    const u4 n1 = ld_u4(&node->n1);
    const u4 n2 = ld_u4(&node->n2);
    const u4 n3 = ld_u4(&node->n3);
    const u4 n4 = ld_u4(&node->n4);
    const float px = bits_f(n0.x), py = bits_f(n0.y), pz = bits_f(n0.z);
    const uint32_t ew = n0.w;
    const float sx = bits_f((ew & 0xffu) << 23) * idx;
    const float sy = bits_f(((ew >> 8) & 0xffu) << 23) * idy;
    const float sz = bits_f(((ew >> 16) & 0xffu) << 23) * idz;
    const float ox = (px - o.x) * idx, oy = (py - o.y) * idy, oz = (pz - o.z) * idz;
    // conservative slack: covers the rounding of (p-o)*idir, of the per-child fma and of the magic-number shift
    const float eps = 4.76837158203125e-07f;   // 2^-21
    // = eps * (|o| + 256 |s|) + |s| / 256, with the two |s| terms folded into one constant (exact: 2^-8 + 2^-13)
    const float spad = 0.0040283203125f;
    const float padx = eps * fabsf(ox) + spad * fabsf(sx);
    const float pady = eps * fabsf(oy) + spad * fabsf(sy);
    const float padz = eps * fabsf(oz) + spad * fabsf(sz);
    const float oxn = (ox - padx) - CRT_BYTE_MAGIC * sx, oxf = (ox + padx) - CRT_BYTE_MAGIC * sx;
    const float oyn = (oy - pady) - CRT_BYTE_MAGIC * sy, oyf = (oy + pady) - CRT_BYTE_MAGIC * sy;
    const float ozn = (oz - padz) - CRT_BYTE_MAGIC * sz, ozf = (oz + padz) - CRT_BYTE_MAGIC * sz;

    uint32_t hitmask = 0u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int half = 0; half < 2; half++) {
        const uint32_t meta4 = half ? n1.w : n1.z;
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = ((is_inner4 >> 4) & 0x01010101u) * 0xffu;
        const uint32_t bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1f1f1f1fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlox = half ? n2.y : n2.x, qloy = half ? n2.w : n2.z, qloz = half ? n3.y : n3.x;
        const uint32_t qhix = half ? n3.w : n3.z, qhiy = half ? n4.y : n4.x, qhiz = half ? n4.w : n4.z;
        const uint32_t nearx = negx ? qhix : qlox, farx = negx ? qlox : qhix;
        const uint32_t neary = negy ? qhiy : qloy, fary = negy ? qloy : qhiy;
        const uint32_t nearz = negz ? qhiz : qloz, farz = negz ? qloz : qhiz;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int j = 0; j < 4; j++) {
            const float t0x = byte_magic(nearx, j, mg) * sx + oxn, t1x = byte_magic(farx, j, mg) * sx + oxf;
            const float t0y = byte_magic(neary, j, mg) * sy + oyn, t1y = byte_magic(fary, j, mg) * sy + oyf;
            const float t0z = byte_magic(nearz, j, mg) * sz + ozn, t1z = byte_magic(farz, j, mg) * sz + ozf;
            const float cmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
            const float cmax = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            if (cmin <= cmax) {
                const uint32_t cb = (child_bits4 >> (8 * j)) & 0xffu;
                const uint32_t bi = (bit_index4 >> (8 * j)) & 0xffu;
                hitmask |= cb << bi;
            }
        }
    }
    ng_x = n1.x;
    ng_y = (hitmask & 0xff000000u) | (ew >> 24);
    tg_x = n1.y;
    tg_y = hitmask & 0x00ffffffu;
}

// Index (relative to the group's first child) of the next node of a node group; clears its hit bit.
CRT_HD uint32_t next_child(uint32_t& ng_y, uint32_t octinv) {
    const int bit = hi_bit(ng_y);
    ng_y &= ~(1u << bit);
    const uint32_t slot = (uint32_t)(bit - 24) ^ octinv;
    return (uint32_t)pop_count(ng_y & ((1u << slot) - 1u) & 0xffu);
}

// ---- resumable traversal ---------------------------------------------------------------------------
// The closest-hit query as a state machine, so that a warp can interleave lanes that are at
// different points of their traversals and replace a finished ray at once (persistent threads with
// dynamic fetch, crt_wavefront.cu) instead of idling until the warp's longest ray is done.
//   begin(o, d, tmin, tmax)  set up a ray
//   step<COUNT>(sc, ctr)     one wide-node visit (8 child boxes) plus the primitive tests it produced;
//                            returns true when the traversal is finished (rec holds the result)
// Entries of the traversal stack are either node groups (hit bits in [31:24]) or primitive groups
// (bits [23:0]); a primitive group that only few lanes of the warp hold is postponed to the stack so
// that the lanes of a warp run the primitive tests together (after Ylitie, Karras, Laine 2017, sec. 4).
#ifndef CRT_POSTPONE_LANES
#define CRT_POSTPONE_LANES 6
#endif

struct StackEntry {
    uint32_t x, y;
};
struct TravStack {   // kept apart from Traverser so that the scalar state stays in registers; all entries in local memory
    uint32_t x[CRT_STACK_ENTRIES], y[CRT_STACK_ENTRIES];
    CRT_HD void set(int k, uint32_t a, uint32_t b) { x[k] = a; y[k] = b; }
    CRT_HD StackEntry get(int k) const { StackEntry e; e.x = x[k]; e.y = y[k]; return e; }
};
#if defined(__CUDACC__)
// The short stack of the kernels: the first SHARED entries of a thread live in shared memory, laid out
// [entry][thread] (8-byte accesses of consecutive threads: conflict free), deeper entries spill to local memory.
// `slice` points at the thread's column of a __shared__ uint2 [SHARED][threads] array.
template <int SHARED, int THREADS>
struct TravStackShared {
    uint2* slice;
    uint32_t lx[CRT_STACK_ENTRIES - SHARED], ly[CRT_STACK_ENTRIES - SHARED];
    __device__ __forceinline__ void set(int k, uint32_t a, uint32_t b) {
        if (k < SHARED) slice[k * THREADS] = make_uint2(a, b);
        else { lx[k - SHARED] = a; ly[k - SHARED] = b; }
    }
    __device__ __forceinline__ StackEntry get(int k) const {
        StackEntry e;
        if (k < SHARED) { const uint2 v = slice[k * THREADS]; e.x = v.x; e.y = v.y; }
        else { e.x = lx[k - SHARED]; e.y = ly[k - SHARED]; }
        return e;
    }
};
#endif

struct Traverser {
    vec3 o, d;
    float idx, idy, idz;
    float tmin;
    HitRec rec;
    uint32_t ng_x, ng_y;   // node group: (first child index, hits[31:24] | imask[7:0])
    uint32_t octinv;
    uint32_t mg;           // byte_magic's constant, see load_magic
    int sp;

    CRT_HD void begin(vec3 o_, vec3 d_, float tmin_, float tmax_) {
        o = o_; d = d_; tmin = tmin_;
        mg = load_magic();
        rec.t = tmax_; rec.prim = 0u; rec.id = -1;
        // reciprocal direction with zero components nudged (box tests only; primitive tests use d)
        const float tiny = 1e-20f;
        const float ddx = fabsf(d.x) < tiny ? (d.x < 0.f ? -tiny : tiny) : d.x;
        const float ddy = fabsf(d.y) < tiny ? (d.y < 0.f ? -tiny : tiny) : d.y;
        const float ddz = fabsf(d.z) < tiny ? (d.z < 0.f ? -tiny : tiny) : d.z;
        idx = f_rcp(ddx); idy = f_rcp(ddy); idz = f_rcp(ddz);
        octinv = (ddx < 0.f ? 0u : 4u) | (ddy < 0.f ? 0u : 2u) | (ddz < 0.f ? 0u : 1u);
        sp = 0;
        ng_x = 0u; ng_y = 0x80000000u;
    }

    template <bool COUNT, class Stack>
    CRT_HD bool step(const SceneView& sc, Stack& st, TraceCounters* ctr, int postpone_lanes = CRT_POSTPONE_LANES) {
        uint32_t tg_x, tg_y;                  // primitive group: (first prim index, hits[23:0])
        if (ng_y & 0xff000000u) {
            const uint32_t base = ng_x;
            uint32_t rest = ng_y;
            const uint32_t rel = next_child(rest, octinv);
            if (rest & 0xff000000u) { st.set(sp, base, rest); sp++; }
            if (COUNT) ctr->nodes++;
#ifdef CRT_NODE_HOOK
            CRT_NODE_HOOK(base + rel);
#endif
            visit_node(sc.nodes + (base + rel), o, idx, idy, idz, tmin, rec.t, octinv, ng_x, ng_y, tg_x, tg_y, mg);
        } else {
            tg_x = ng_x;
            tg_y = ng_y;
            ng_x = 0u;
            ng_y = 0u;
        }

        while (tg_y) {
#if defined(__CUDA_ARCH__)
            // few lanes hold primitives and this lane still has nodes to visit: keep the group for later
            if ((ng_y & 0xff000000u) && __popc(__activemask()) < postpone_lanes) {
                st.set(sp, tg_x, tg_y); sp++;
                break;
            }
#endif
            // two records per trip: both loads in flight before the first test needs its data
            const int i0 = hi_bit(tg_y);
            tg_y &= ~(1u << i0);
            const bool two = tg_y != 0u;
            const int i1 = two ? hi_bit(tg_y) : i0;
            tg_y &= ~(1u << i1);
            const PrimLoad r0 = load_prim(sc.prims, tg_x + (uint32_t)i0);
            const PrimLoad r1 = load_prim(sc.prims, tg_x + (uint32_t)i1);
            if (COUNT) ctr->prims += two ? 2u : 1u;
#ifdef CRT_PRIM_HOOK
            CRT_PRIM_HOOK(tg_x + (uint32_t)i0);
            if (two) CRT_PRIM_HOOK(tg_x + (uint32_t)i1);
#endif
            test_prim_loaded(r0, tg_x + (uint32_t)i0, o, d, tmin, rec);
            if (two) test_prim_loaded(r1, tg_x + (uint32_t)i1, o, d, tmin, rec);
        }

        if ((ng_y & 0xff000000u) == 0u) {
            if (sp == 0) return true;
            sp--;
            const StackEntry e = st.get(sp);
            ng_x = e.x;
            ng_y = e.y;
        }
        return false;
    }
};

// Primitives kept out of the BVH (crt_bvh_build.h "giant primitives"): tested by every ray before it traverses.
template <bool COUNT>
CRT_HD void trace_giants(const SceneView& sc, vec3 o, vec3 d, float tmin, HitRec& rec, TraceCounters* ctr) {
    for (uint32_t g = 0; g < sc.giant_count; g++) {
        if (COUNT) ctr->prims++;
        test_prim(sc.prims, sc.giant_base + g, o, d, tmin, rec);
    }
}

// Closest hit in [tmin, tmax]. COUNT switches the work counters on (crt_render_counted).
template <bool COUNT, class Stack>
CRT_HD bool trace_closest_on(Stack& st, const SceneView& sc, vec3 o, vec3 d, float tmin, float tmax, HitRec& rec, TraceCounters* ctr) {
    Traverser tr;
    tr.begin(o, d, tmin, tmax);
    trace_giants<COUNT>(sc, o, d, tmin, tr.rec, ctr);
    while (!tr.step<COUNT>(sc, st, ctr)) {}
    rec = tr.rec;
    return rec.id >= 0;
}
template <bool COUNT>
CRT_HD bool trace_closest(const SceneView& sc, vec3 o, vec3 d, float tmin, float tmax, HitRec& rec, TraceCounters* ctr) {
    TravStack st;
    return trace_closest_on<COUNT>(st, sc, o, d, tmin, tmax, rec, ctr);
}

}  // namespace crt
