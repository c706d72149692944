// crt_wavefront.cu — wavefront path tracer: persistent-thread begin / extend / shade kernels over
// SoA path state with ballot+popc queue compaction (sm_100a).
//
// The reference traces all samples of a pixel inside one thread (reference:
// CudaRayTracer/src/CUDAKernels.h:147-166). Its per-pixel cuRAND stream makes the samples of a
// pixel sequentially dependent (sample s+1 starts where sample s stopped drawing), so the unit of
// parallelism here is still the pixel: every pixel owns ONE path at a time, and the wavefront is
// "one ray per unfinished pixel". Each iteration runs
//     shade  : consume the previous hit (scatter / emission / sky), finish or continue the path,
//              regenerate a camera ray when the path ended, apply Russian roulette for the next
//              bounce, and compact the surviving pixels into the next queue
//     extend : closest-hit query for every queued ray (the lean, traversal-only kernel)
// Draw order per pixel is exactly the megakernel's, so both modes produce identical bits.
//
// Path state (one entry per pixel):
//   A = (o.x, o.y, o.z, thr.x)   B = (d.x, d.y, d.z, thr.y)   C = (t, prim, thr.z, book)
//   book = sample index (bits 0-23) | bounce (bits 24-31)
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "crt_internal.h"
#include "crt_shade.h"

namespace crt {

#ifndef CRT_FETCH_LANES
#define CRT_FETCH_LANES 12
#endif

// The pixel set is split into CRT_WF_PARTS interleaved parts (2048-pixel tiles, round robin), each with its own
// queue and its own extend -> shade chain on its own stream: while one part's extend kernel drains (its last,
// longest rays run alone and latency bound: 16 % of the kernel's duration had idle SMs in ncu), the other
// part's kernels fill the machine.
#define CRT_WF_MAX_PARTS 4
struct WfPart {
    uint32_t* queue[2] = {nullptr, nullptr};
    uint32_t* qcount = nullptr;     // [0],[1] queue sizes, [2] fetch cursor of extend
    uint32_t* tile_count = nullptr; // per-tile survivor counts / offsets of the compaction
    uint32_t* h_qcount = nullptr;   // pinned
    uint2* suspended[2] = {nullptr, nullptr};   // suspended traversals, [launch parity][extend thread][CRT_SUSPEND_WORDS]
    uint32_t extend_launches = 0;   // parity of the next extend launch
    uint32_t cap = 0;               // pixels of this part
    uint32_t live = 0;              // queue length of the current batch (host copy)
    int cur = 0;
    bool done = false;
    int tail = 0;                   // 0 = wavefront iterations, 1 = handed to k_wf_tail (closing extend still to run), 2 = tail stages
    cudaStream_t stream = nullptr;  // part 0 runs on the caller's stream
    cudaEvent_t ev_done = nullptr;
    // CRT_EXTEND_OCTANT experiment: a copy of the queue sorted by ray-direction octant, read by extend only
    uint32_t* oct_keys[2] = {nullptr, nullptr};
    uint32_t* oct_vals[2] = {nullptr, nullptr};
    void* oct_tmp = nullptr;
    size_t oct_tmp_bytes = 0;
};

// One batch of iterations (16 x parts x (extend, shade) + the compaction triples + the queue-length read-backs) as an
// instantiated CUDA graph, keyed by everything its launches depend on (kernel arguments included).
struct WfGraph {
    std::vector<unsigned char> sig;
    cudaGraphExec_t exec = nullptr;
    int seen = 0;
};

struct WavefrontState {
    std::vector<WfGraph> graphs;
    int batch = 16;                  // env CRT_WF_BATCH (1..16): iterations between two queue compactions / queue-length read-backs
    int use_graph = 1;               // env CRT_WF_GRAPH: replay a batch whose launches were seen before as one CUDA graph launch
    int extend_octant = 0;           // env CRT_EXTEND_OCTANT: bracket experiment, extend reads a copy of the queue sorted by direction octant (cub radix sort per iteration)
    int blocked = 0;                 // env CRT_WF_BLOCKED: temporal blocking experiment, a part runs its whole batch before the next part starts (one stream)
    float4* A = nullptr;
    float4* B = nullptr;
    float4* C = nullptr;
    WfPart part[CRT_WF_MAX_PARTS];
    int parts = 2;   // env CRT_WF_PARTS
    cudaStream_t side_stream[CRT_WF_MAX_PARTS] = {};   // [p] for part p >= 1
    cudaEvent_t ev_fork = nullptr;
    size_t n = 0;
    int blocks_extend = 0, blocks_shade = 0;
    int shade_bulk = 0;              // env CRT_SHADE_BULK=1: shade's path state arrives through cp.async.bulk (k_wf_shade_bulk); 0 = per-thread loads (k_wf_shade). Measured equal
    int extend_reverse = 1;          // env CRT_EXTEND_REVERSE: extend consumes the queue back to front (L2 reuse of what shade just wrote)
    int fetch_lanes = CRT_FETCH_LANES, postpone_lanes = CRT_POSTPONE_LANES, coop = 0, shade_persistent = 1;   // tuning knobs (env CRT_FETCH_LANES / CRT_POSTPONE_LANES)
    int drain_rounds = 4;            // env CRT_DRAIN_ROUNDS: rounds a warp keeps going after the queue ran out, then it suspends its rays (0 = never suspend)
    long long tail_live = 150000;    // env CRT_TAIL_LIVE: a part whose queue is at most this long is finished by k_wf_tail stages (0 = never)
    int tail_iters = 64;             // env CRT_TAIL_ITERS: path segments per pixel and tail stage
    int tail_lanes_min = 1;          // env CRT_TAIL_LANES_MIN: lanes of a warp that take a pixel, at least
    int tail_blocks = 4;             // resident k_wf_tail CTAs per SM (occupancy query)
    int tail_shade_lanes = 8;        // env CRT_TAIL_SHADE_LANES: of 32; lanes that have to be waiting before a tail warp runs its shade step
    long long drain_min_live = -1;   // env CRT_DRAIN_MIN_LIVE: suspend only while the part's queue is at least this long (-1 = one ray per extend thread)
    cudaEvent_t ev[3 * 16 * CRT_WF_MAX_PARTS + 2 + 4 * CRT_WF_MAX_PARTS] = {};   // CRT_RENDER_TIME_KERNELS: 3 events per (iteration, part) of a batch, begin pair, tail pair and compaction pair per part
};

struct WfView {
    float4* A;
    float4* B;
    float4* C;
};

__device__ __forceinline__ Rng wf_load_rng(const uint32_t* planes, size_t n, size_t pix) {
    Rng r;
    r.v0 = planes[pix]; r.v1 = planes[n + pix]; r.v2 = planes[2 * n + pix];
    r.v3 = planes[3 * n + pix]; r.v4 = planes[4 * n + pix]; r.d = planes[5 * n + pix];
    return r;
}
__device__ __forceinline__ void wf_store_rng(uint32_t* planes, size_t n, size_t pix, const Rng& r) {
    planes[pix] = r.v0; planes[n + pix] = r.v1; planes[2 * n + pix] = r.v2;
    planes[3 * n + pix] = r.v3; planes[4 * n + pix] = r.v4; planes[5 * n + pix] = r.d;
}

// ---- queue compaction -----------------------------------------------------------------------------
// A pixel leaves the wavefront only when all its samples are done, so the queue shrinks slowly. shade marks
// finished entries QUEUE_DEAD in place and every CRT_WF_BATCH iterations the queue is compacted with an
// ORDER-PRESERVING ballot/popc scan: the queue stays sorted by pixel index, so the path-state and RNG-plane
// accesses of extend and shade stay coalesced for the whole frame. (Appending survivors with one atomic per
// warp, the first version, shuffles the queue a little at every departure; by iteration ~3000 of a 2000-spp
// frame neighbouring entries came from unrelated pixels and shade ran 3x slower.)
#define QUEUE_DEAD 0xffffffffu
// A ray that extend parks (see k_wf_extend, "drain") is marked in its pixel's path state, twice:
//  * C.x = PARKED_T instead of a hit distance: shade skips the pixel. Overwritten when the ray finishes.
//  * the sign bit of A.w (throughput.x, never negative otherwise): the ray is served from the parked-ray list,
//    so the queue fetch of extend skips the pixel. It has to outlive the ray (the list entry may be finished
//    before another warp meets the pixel's queue entry), so it stays until shade rewrites A.
// PARKED_LANE tags a lane's `pix` register while it runs a resumed ray.
#define PARKED_T (-2.0f)
#define PARKED_SIGN 0x80000000u
#define PARKED_LANE 0x80000000u
#define PIXEL_MASK 0x7fffffffu
// One parked traversal: [0] = (pixel, stack depth), [1] = (t bits, winning primitive record), [2] = (winning id, -),
// [3] = current node group, [4..] = the traversal stack. At most one per extend thread and launch; two lists per
// part, written and read alternately (qcount[4 + parity] = length).
#define CRT_SUSPEND_WORDS (4 + CRT_STACK_ENTRIES)
#define CRT_COMPACT_TILE 2048   // entries per CTA (256 threads x 8)

__global__ void __launch_bounds__(256) k_compact_count(const uint32_t* queue, const uint32_t* qcount, int cur, uint32_t* tile_count) {
    const uint32_t count = qcount[cur];
    const uint32_t tile0 = blockIdx.x * CRT_COMPACT_TILE;
    if (tile0 >= count) { if (threadIdx.x == 0) tile_count[blockIdx.x] = 0u; return; }
    uint32_t n = 0;
#pragma unroll
    for (int k = 0; k < CRT_COMPACT_TILE / 256; k++) {
        const uint32_t i = tile0 + k * 256 + threadIdx.x;
        n += (i < count && queue[i] != QUEUE_DEAD) ? 1u : 0u;
    }
    __shared__ uint32_t warp_n[8];
    n = __reduce_add_sync(0xffffffffu, n);
    if ((threadIdx.x & 31) == 0) warp_n[threadIdx.x >> 5] = n;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int k = 0; k < 8; k++) t += warp_n[k];
        tile_count[blockIdx.x] = t;
    }
}

// one CTA: exclusive scan of the tile counts in place, total -> qcount[cur ^ 1]
__global__ void __launch_bounds__(1024) k_compact_scan(uint32_t* tile_count, int tiles, uint32_t* qcount, int cur) {
    __shared__ uint32_t warp_sum[32];
    __shared__ uint32_t carry_s;
    if (threadIdx.x == 0) carry_s = 0u;
    __syncthreads();
    for (int base = 0; base < tiles; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < tiles ? tile_count[i] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, incl, d);
            if ((threadIdx.x & 31) >= d) incl += u;
        }
        if ((threadIdx.x & 31) == 31) warp_sum[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            uint32_t w = warp_sum[threadIdx.x];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, w, d);
                if (threadIdx.x >= d) w += u;
            }
            warp_sum[threadIdx.x] = w;
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        const uint32_t before = (threadIdx.x >> 5) ? warp_sum[(threadIdx.x >> 5) - 1] : 0u;
        if (i < tiles) tile_count[i] = carry + before + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + before + incl;
        __syncthreads();
    }
    if (threadIdx.x == 0) { qcount[cur ^ 1] = carry_s; qcount[2] = 0u; }
}

__global__ void __launch_bounds__(256) k_compact_scatter(const uint32_t* queue_in, uint32_t* queue_out, const uint32_t* qcount, int cur,
                                                         const uint32_t* tile_offset) {
    const uint32_t count = qcount[cur];
    const uint32_t tile0 = blockIdx.x * CRT_COMPACT_TILE;
    if (tile0 >= count) return;
    __shared__ uint32_t warp_n[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t out = tile_offset[blockIdx.x];
    for (int k = 0; k < CRT_COMPACT_TILE / 256; k++) {
        const uint32_t i = tile0 + k * 256 + threadIdx.x;
        const uint32_t v = i < count ? queue_in[i] : QUEUE_DEAD;
        const bool keep = v != QUEUE_DEAD;
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_n[warp] = (uint32_t)__popc(m);
        __syncthreads();
        uint32_t before = 0, all = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) { const uint32_t c = warp_n[q]; if (q < warp) before += c; all += c; }
        if (keep) queue_out[out + before + (uint32_t)__popc(m & ((1u << lane) - 1u))] = v;
        out += all;
        __syncthreads();
    }
}

// ---- CRT_EXTEND_OCTANT (experiment): sort key = direction octant of the queued ray, 8 = finished entry ----------
__global__ void __launch_bounds__(256) k_octant_keys(const uint32_t* queue, const uint32_t* qcount, int cur, const float4* B, uint32_t* keys, uint32_t* vals, uint32_t n) {
    const uint32_t count = qcount[cur];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t pix = i < count ? queue[i] : QUEUE_DEAD;
        uint32_t key = 8u;
        if (pix != QUEUE_DEAD) {
            const float4 b = B[pix];
            key = (b.x < 0.f ? 4u : 0u) | (b.y < 0.f ? 2u : 0u) | (b.z < 0.f ? 1u : 0u);
        }
        keys[i] = key;
        vals[i] = pix;
    }
}

// ---- begin: first camera ray of every pixel, queue = all pixels ---------------------------------
struct BeginQueues {
    uint32_t* queue[CRT_WF_MAX_PARTS];
    uint32_t* qcount[CRT_WF_MAX_PARTS];
    uint32_t cap[CRT_WF_MAX_PARTS];
    int parts;
};
__global__ void __launch_bounds__(256) k_wf_begin(RenderArgs a, WfView w, BeginQueues q) {
    const size_t n = (size_t)a.rc.width * a.rc.height;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const int x = (int)(i % (size_t)a.rc.width), y = (int)(i / (size_t)a.rc.width);
        Rng rng = wf_load_rng(a.fb.rng, n, i);
        vec3 o, d;
        camera_ray(a.cam, x, y, a.rc.width, a.rc.height, rng, o, d);
        wf_store_rng(a.fb.rng, n, i, rng);
        w.A[i] = make_float4(o.x, o.y, o.z, 1.f);
        w.B[i] = make_float4(d.x, d.y, d.z, 1.f);
        w.C[i] = make_float4(0.f, 0.f, 1.f, __uint_as_float(0u));
        if (!(a.flags & CRT_RENDER_ACCUMULATE)) a.fb.accum[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        // tile t of CRT_COMPACT_TILE pixels belongs to part t % parts; only the last tile can be partial
        const size_t tile = i / CRT_COMPACT_TILE;
        const int p = (int)(tile % (size_t)q.parts);
        q.queue[p][(tile / (size_t)q.parts) * CRT_COMPACT_TILE + (i % CRT_COMPACT_TILE)] = (uint32_t)i;
    }
    if (blockIdx.x == 0 && threadIdx.x < q.parts) {
        uint32_t* c = q.qcount[threadIdx.x];
        c[0] = q.cap[threadIdx.x]; c[1] = 0u; c[2] = 0u; c[4] = 0u; c[5] = 0u;
    }
}

// ---- extend: closest hit for every queued ray -------------------------------------------------------
// Persistent warps, two ideas against SIMT divergence (ncu on the one-ray-per-thread version showed 8 of
// 32 lanes active per instruction):
//  1. dynamic fetch: a warp keeps up to 32 rays in flight; when `fetch_lanes` lanes have finished they
//     take the next rays from the queue together (one atomic per refill), so a long ray (one that wanders
//     through the mesh) does not hold 31 finished lanes idle. qcount[2] is the shared fetch cursor.
//  2. warp-cooperative primitive tests: a node visit leaves each lane with 0..n primitives to test, so a
//     per-lane loop runs with ~6 lanes. Instead the (ray, primitive) pairs of the whole warp are pooled in
//     shared memory and dealt out one pair per lane; a hit updates the owning ray's closest hit with a
//     64-bit shared-memory atomicMin on (t bits << 32 | ~id), which is exactly the order-independent
//     acceptance rule of test_prim (smaller t, then larger primitive id).
struct WarpPool {
    float ray[6][32];                 // o.xyz, d.xyz of the lane's ray
    unsigned long long best[32];      // (t bits << 32) | (0xffffffff - id)
    uint32_t win[32];                 // index of the winning primitive record (0xffffffff = none)
    uint32_t list[256];               // (owner lane << 27) | primitive record index
};

#ifndef CRT_EXTEND_THREADS
#define CRT_EXTEND_THREADS 256
#endif
#ifndef CRT_SHADE_PREFETCH
#define CRT_SHADE_PREFETCH 1   // 0: k_wf_shade without the L2 prefetches of the next entry's state (experiment)
#endif
#ifndef CRT_PRIM_PIPE
#define CRT_PRIM_PIPE 0   // 1: one primitive record per trip of the test loop, software pipelined (experiment)
#endif
#ifndef CRT_PREFETCH_NEXT
#define CRT_PREFETCH_NEXT 0   // 1: L2 prefetch of the node on top of the traversal stack during every node visit (experiment)
#endif
#ifndef CRT_SMEM_STACK
#define CRT_SMEM_STACK 4   // entries of the traversal stack kept in shared memory (0 = all in local memory); see DESIGN.md 3.2
#endif
// Path state and queue words are touched once per launch: their loads are ld.global.cg (L2 only), so that they do not
// push nodes and primitive records out of L1 (0: plain loads, 1: extend only, 2: extend and shade; +0.3 % on both scenes,
// DESIGN.md 3.4)
#ifndef CRT_STATE_BYPASS_L1
#define CRT_STATE_BYPASS_L1 2
#endif
#if CRT_STATE_BYPASS_L1 >= 1
#define CRT_LD_STATE_E(p) __ldcg(p)
#else
#define CRT_LD_STATE_E(p) (*(p))
#endif
#if CRT_STATE_BYPASS_L1 >= 2
#define CRT_LD_STATE_S(p) __ldcg(p)
#else
#define CRT_LD_STATE_S(p) (*(p))
#endif
#ifndef CRT_SHADE_PREFETCH_PRIM
#define CRT_SHADE_PREFETCH_PRIM 0   // 1: shade reads the next entry's hit early and prefetches its primitive record into L1 at the end of the trip (experiment)
#endif
#ifndef CRT_STATE_STORE_CG
#define CRT_STATE_STORE_CG 0        // 1: path-state stores of shade as st.global.cg (experiment)
#endif
#ifndef CRT_DEBUG_WATCHDOG
#define CRT_DEBUG_WATCHDOG 0   // 1 (debug builds only): a warp whose loop passes 2^22 trips prints its lanes' state and leaves
#endif
#ifndef CRT_EXTEND_MIN_BLOCKS
#define CRT_EXTEND_MIN_BLOCKS (1024 / CRT_EXTEND_THREADS)
#endif
template <bool COUNT, bool COOP>
__global__ void __launch_bounds__(CRT_EXTEND_THREADS, CRT_EXTEND_MIN_BLOCKS) k_wf_extend(SceneView scene, WfView w, uint32_t* queue, uint32_t* qcount, int cur,
                                                   unsigned long long* counters, int fetch_lanes, int drain_rounds,
                                                   uint2* susp_out, const uint2* susp_in, int par, int rev) {
    __shared__ WarpPool pools[COOP ? CRT_EXTEND_THREADS / 32 : 1];
    WarpPool& S = pools[COOP ? threadIdx.x >> 5 : 0];
    // work items: first the rays the previous launch parked (they are the long ones: started first, they have the
    // whole launch to finish), then the queue
    const uint32_t nres = qcount[4 + (par ^ 1)];
    const uint32_t count = nres + qcount[cur];
    const int lane = threadIdx.x & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const float tmin = 0.001f;
    const uint32_t magic = load_magic();   // byte_magic's constant, kept in a register for the whole launch
    uint32_t c_nodes = 0, c_prims = 0, c_rays = 0;
    // Traversal stack: the first CRT_SMEM_STACK entries live in shared memory ([entry][thread]: conflict-free 8-byte
    // accesses), deeper ones spill to the thread's local-memory array. A ray of the headline scene makes 2.5 node
    // visits, so its stack seldom passes 3 entries; the deep entries are for the 5 M-triangle tree.
#if CRT_SMEM_STACK > 0
    __shared__ uint2 sstack[CRT_SMEM_STACK][CRT_EXTEND_THREADS];
    uint32_t lx[CRT_STACK_ENTRIES - CRT_SMEM_STACK], ly[CRT_STACK_ENTRIES - CRT_SMEM_STACK];
    auto st_set = [&](int k, uint32_t x, uint32_t y) {
        if (k < CRT_SMEM_STACK) sstack[k][threadIdx.x] = make_uint2(x, y);
        else { lx[k - CRT_SMEM_STACK] = x; ly[k - CRT_SMEM_STACK] = y; }
    };
    auto st_get = [&](int k) -> uint2 {
        if (k < CRT_SMEM_STACK) return sstack[k][threadIdx.x];
        return make_uint2(lx[k - CRT_SMEM_STACK], ly[k - CRT_SMEM_STACK]);
    };
#else
    TravStack st;
    auto st_set = [&](int k, uint32_t x, uint32_t y) { st.x[k] = x; st.y[k] = y; };
    auto st_get = [&](int k) -> uint2 { return make_uint2(st.x[k], st.y[k]); };
#endif
    vec3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f);
    HitRec rec;
    rec.t = INFINITY; rec.prim = 0u; rec.id = -1;
    float idx = 1.f, idy = 1.f, idz = 1.f;
    uint32_t octinv = 0, ng_x = 0, ng_y = 0;
    int sp = 0;
    bool has_ray = false, exhausted = false;
    uint32_t pix = 0;   // | PARKED_LANE while the lane runs a resumed ray
    int drain = drain_rounds;
#if CRT_DEBUG_WATCHDOG
    uint32_t wd_trips = 0;
#endif
    while (true) {
#if CRT_DEBUG_WATCHDOG
        if (++wd_trips > (1u << 22)) {
            printf("WATCHDOG k_wf_extend blk %d thr %d has_ray %d exhausted %d pix %08x sp %d ng %08x %08x rec.t %g o %g %g %g d %g %g %g count %u nres %u cursor %u drain_rounds %d\n",
                   blockIdx.x, threadIdx.x, (int)has_ray, (int)exhausted, pix, sp, ng_x, ng_y, rec.t, o.x, o.y, o.z, d.x, d.y, d.z, count, nres, qcount[2], drain_rounds);
            break;
        }
#endif
        unsigned idle = __ballot_sync(0xffffffffu, !has_ray);
        if (!exhausted && __popc(idle) >= fetch_lanes) {
            const int n = __popc(idle), leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if (lane == leader) base = atomicAdd(&qcount[2], (uint32_t)n);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (base + (uint32_t)n >= count) exhausted = true;
            const uint32_t i = base + (uint32_t)__popc(idle & lt_mask);
            const uint2* r = susp_in + (size_t)i * CRT_SUSPEND_WORDS;
            bool take = !has_ray && i < count;
            if (take) {
                // the queue is consumed back to front: shade, which ran just before on this part, swept it front to back,
                // so the path states it wrote last (still in L2) are the ones read first; the next shade starts at the
                // front, where this kernel ends
                pix = i < nres ? (r[0].x | PARKED_LANE) : CRT_LD_STATE_E(&queue[rev ? count - 1u - i : i - nres]);
                take = pix != QUEUE_DEAD;
            }
            float4 A = make_float4(0.f, 0.f, 0.f, 0.f), B = A;
            if (take) {
                A = CRT_LD_STATE_E(&w.A[pix & PIXEL_MASK]);
                B = CRT_LD_STATE_E(&w.B[pix & PIXEL_MASK]);
                // a parked pixel met in the queue: its ray is in the list, not here
                take = (pix & PARKED_LANE) || !(__float_as_uint(A.w) & PARKED_SIGN);
            }
            if (take) {
                o = mk3(A.x, A.y, A.z);
                d = mk3(B.x, B.y, B.z);
                rec.t = INFINITY; rec.prim = 0u; rec.id = -1;
                const float tiny = 1e-20f;   // zero direction components nudged for the box tests only
                const float ddx = fabsf(B.x) < tiny ? (B.x < 0.f ? -tiny : tiny) : B.x;
                const float ddy = fabsf(B.y) < tiny ? (B.y < 0.f ? -tiny : tiny) : B.y;
                const float ddz = fabsf(B.z) < tiny ? (B.z < 0.f ? -tiny : tiny) : B.z;
                idx = f_rcp(ddx); idy = f_rcp(ddy); idz = f_rcp(ddz);   // = 1.0f / x (both correctly rounded), without the division's slow path
                octinv = (ddx < 0.f ? 0u : 4u) | (ddy < 0.f ? 0u : 2u) | (ddz < 0.f ? 0u : 1u);
                if (pix & PARKED_LANE) {
                    // a traversal parked by the previous extend launch: closest hit so far, node group and stack come back
                    const uint2 r0 = r[0], r1 = r[1], r2 = r[2], r3 = r[3];
                    rec.t = __uint_as_float(r1.x); rec.prim = r1.y; rec.id = (int32_t)r2.x;
                    sp = (int)r0.y; ng_x = r3.x; ng_y = r3.y;
                    for (int k = 0; k < sp; k++) { const uint2 e = r[4 + k]; st_set(k, e.x, e.y); }
                } else {
                sp = 0; ng_x = 0u; ng_y = 0x80000000u;
                if (COOP) {
                    S.ray[0][lane] = A.x; S.ray[1][lane] = A.y; S.ray[2][lane] = A.z;
                    S.ray[3][lane] = B.x; S.ray[4][lane] = B.y; S.ray[5][lane] = B.z;
                    S.best[lane] = ((unsigned long long)0x7f800000u << 32) | 0xffffffffull;   // t = +inf
                    S.win[lane] = 0xffffffffu;
                }
                // primitives outside the BVH first: they bound the interval the traversal has to search
                for (uint32_t g = 0; g < scene.giant_count; g++) {
                    if (COUNT) c_prims++;
                    test_prim(scene.prims, scene.giant_base + g, o, d, tmin, rec);
                }
                if (COOP && rec.id >= 0) {
                    S.best[lane] = ((unsigned long long)__float_as_uint(rec.t) << 32) | (unsigned long long)(0xffffffffu - (uint32_t)rec.id);
                    S.win[lane] = rec.prim;
                }
                c_rays++;
                }
                has_ray = true;
            }
            idle = __ballot_sync(0xffffffffu, !has_ray);
        }
        if (idle == 0xffffffffu) {
            if (exhausted) break;
            continue;   // a refill that met only finished (dead) queue entries: fetch again
        }
        // drain: once the queue has run out a warp goes on for `drain_rounds` rounds, then parks what is left.
        // The last rays of a launch are its longest (tens of node visits, each a dependent chain of loads)
        // and would hold the kernel open, latency bound, with most SMs idle: measured 0.13 ms of a 0.27 ms
        // launch. A parked ray's pixel is marked (PARKED_SIGN), shade skips it, and the next extend launch starts
        // with the parked rays, each taken up where it stopped, alongside a full queue of other rays.
        if (exhausted && drain_rounds > 0 && --drain < 0) {
            if (has_ray) {
                const uint32_t slot = atomicAdd(&qcount[4 + par], 1u);
                uint2* r = susp_out + (size_t)slot * CRT_SUSPEND_WORDS;
                r[0] = make_uint2(pix & PIXEL_MASK, (uint32_t)sp);
                r[1] = make_uint2(__float_as_uint(rec.t), rec.prim);
                r[2] = make_uint2((uint32_t)rec.id, 0u);
                r[3] = make_uint2(ng_x, ng_y);
                for (int k = 0; k < sp; k++) r[4 + k] = st_get(k);
                w.C[pix & PIXEL_MASK].x = PARKED_T;
                if (!(pix & PARKED_LANE)) {
                    uint32_t* aw = reinterpret_cast<uint32_t*>(&w.A[pix].w);
                    *aw |= PARKED_SIGN;
                }
            }
            break;
        }

        // node round: every lane with a ray visits its next wide node (a lane with a ray always has one pending)
        uint32_t tg_x = 0u, tg_y = 0u;
        if (has_ray) {
            const uint32_t base = ng_x;
            uint32_t rest = ng_y;
            const uint32_t rel = next_child(rest, octinv);
            if (rest & 0xff000000u) { st_set(sp, base, rest); sp++; }
#if CRT_PREFETCH_NEXT
            // experiment ("two node visits in flight"): the node this lane visits next unless the current
            // visit finds children of its own -- the first hit child of the group on top of the stack -- is requested
            // into L2 while the current visit computes
            if (sp > 0) {
                const uint2 top = st_get(sp - 1);
                uint32_t tr = top.y;
                const uint32_t r2 = next_child(tr, octinv);
                asm volatile("prefetch.global.L2 [%0];" :: "l"(scene.nodes + (top.x + r2)));
            }
#endif
            if (COUNT) c_nodes++;
            const float tmax = COOP ? __uint_as_float((uint32_t)(S.best[lane] >> 32)) : rec.t;
            visit_node(scene.nodes + (base + rel), o, idx, idy, idz, tmin, tmax, octinv, ng_x, ng_y, tg_x, tg_y, magic);
        }

        __syncwarp();
        if (!COOP) {   // every lane loops over its own primitives; closest hit kept in registers
            // two records per trip: both loads are in flight before the first test needs its data (leaves of
            // the wall quads hold triangle pairs), which halves the exposed L1/L2 latency of this phase
#if CRT_PRIM_PIPE
            // one record per trip, the next record's loads issued before the current one is tested: a warp whose
            // lanes hold 1, 1, 3 primitives makes 3 trips of one test instead of 2 trips of two
            if (tg_y) {
                int i0 = 31 - __clz((int)tg_y);
                tg_y &= ~(1u << i0);
                uint32_t cur_idx = tg_x + (uint32_t)i0;
                PrimLoad cur = load_prim(scene.prims, cur_idx);
                while (true) {
                    const bool more = tg_y != 0u;
                    uint32_t nxt_idx = cur_idx;
                    PrimLoad nxt = cur;
                    if (more) {
                        i0 = 31 - __clz((int)tg_y);
                        tg_y &= ~(1u << i0);
                        nxt_idx = tg_x + (uint32_t)i0;
                        nxt = load_prim(scene.prims, nxt_idx);
                    }
                    if (COUNT) c_prims++;
                    test_prim_loaded(cur, cur_idx, o, d, tmin, rec);
                    if (!more) break;
                    cur = nxt;
                    cur_idx = nxt_idx;
                }
            }
#else
            while (tg_y) {
                const int i0 = 31 - __clz((int)tg_y);
                tg_y &= ~(1u << i0);
                const bool two = tg_y != 0u;
                const int i1 = two ? 31 - __clz((int)tg_y) : i0;
                tg_y &= ~(1u << i1);
                const PrimLoad r0 = load_prim(scene.prims, tg_x + (uint32_t)i0);
                const PrimLoad r1 = load_prim(scene.prims, tg_x + (uint32_t)i1);
                if (COUNT) c_prims += two ? 2u : 1u;
                test_prim_loaded(r0, tg_x + (uint32_t)i0, o, d, tmin, rec);
                if (two) test_prim_loaded(r1, tg_x + (uint32_t)i1, o, d, tmin, rec);
            }
#endif
        } else {
        // cooperative primitive rounds: pool the warp's (ray, primitive) pairs, one pair per lane
        while (true) {
            int cnt = __popc(tg_y);
            if (cnt > 8) cnt = 8;
            int incl = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += v;
            }
            const int total = __shfl_sync(0xffffffffu, incl, 31);
            if (total == 0) break;
            int pos = incl - cnt;
            for (int j = 0; j < cnt; j++) {
                const int i = 31 - __clz((int)tg_y);
                tg_y &= ~(1u << i);
                S.list[pos++] = ((uint32_t)lane << 27) | (tg_x + (uint32_t)i);
            }
            __syncwarp();
            for (int k0 = 0; k0 < total; k0 += 32) {
                const int k = k0 + lane;
                bool hit = false;
                unsigned long long key = 0ull;
                uint32_t owner = 0u, pi = 0u;
                if (k < total) {
                    const uint32_t e = S.list[k];
                    owner = e >> 27;
                    pi = e & 0x07ffffffu;
                    const vec3 ro = mk3(S.ray[0][owner], S.ray[1][owner], S.ray[2][owner]);
                    const vec3 rd = mk3(S.ray[3][owner], S.ray[4][owner], S.ray[5][owner]);
                    const float tmax = __uint_as_float((uint32_t)(S.best[owner] >> 32));
                    float t;
                    int32_t id;
                    if (COUNT) c_prims++;
                    hit = prim_hit(scene.prims, pi, ro, rd, tmin, tmax, t, id);
                    if (hit) {
                        key = ((unsigned long long)__float_as_uint(t) << 32) | (unsigned long long)(0xffffffffu - (uint32_t)id);
                        atomicMin(&S.best[owner], key);
                    }
                }
                __syncwarp();
                if (hit && S.best[owner] == key) S.win[owner] = pi;
                __syncwarp();
            }
        }
        }
        // end of the round: a lane whose node group is used up takes the next one from its stack, or is done
        // (done lanes are refilled at the top of the loop, so they do not sit through another round)
        if (has_ray && (ng_y & 0xff000000u) == 0u) {
            if (sp == 0) {
                float t;
                uint32_t win;
                if (COOP) { win = S.win[lane]; t = __uint_as_float((uint32_t)(S.best[lane] >> 32)); }
                else { win = rec.id >= 0 ? rec.prim : 0xffffffffu; t = rec.t; }
                pix &= PIXEL_MASK;
                float2* c = reinterpret_cast<float2*>(&w.C[pix]);
                *c = make_float2(win != 0xffffffffu ? t : -1.f, __uint_as_float(win));
                has_ray = false;
            } else {
                sp--;
                const uint2 e = st_get(sp);
                ng_x = e.x;
                ng_y = e.y;
            }
        }
    }
    c_rays = __reduce_add_sync(0xffffffffu, c_rays);
    if (lane == 0 && c_rays) atomicAdd(&counters[0], (unsigned long long)c_rays);
    if (COUNT) {
        c_nodes = __reduce_add_sync(0xffffffffu, c_nodes);
        c_prims = __reduce_add_sync(0xffffffffu, c_prims);
        if (lane == 0 && (c_nodes | c_prims)) { atomicAdd(&counters[2], (unsigned long long)c_nodes); atomicAdd(&counters[3], (unsigned long long)c_prims); }
    }
}

// One shade step of a pixel's path, on registers: consume the hit (t < 0 = miss), continue / finish the path,
// start the pixel's next sample when the path ended, draw Russian roulette for the next bounce. Returns false
// when the pixel's last sample is done. Shared by k_wf_shade (one step per launch) and k_wf_tail (looped).
__device__ __forceinline__ bool shade_step(const RenderArgs& a, uint32_t pix, vec3& o, vec3& d, vec3& thr, int& sample, int& bounce, Rng& rng,
                                           float hit_t, uint32_t hit_prim, float4& acc, bool& touched) {
    bool keep;
    bool ended = false;
    vec3 result = mk3(0.f, 0.f, 0.f);
    if (!(hit_t < 0.f)) {                                  // extend stores t = -1 for a miss
        HitRec h;
        h.t = hit_t; h.prim = hit_prim; h.id = 0;
        const Surface s = surface_at(a.scene, h, o, d);
        if (s.mat < a.scene.num_mats) {
            const MatRec m = a.scene.mats[s.mat];
            vec3 att, no, nd;
            if (scatter(m, s, d, rng, att, no, nd)) {
                thr = v_mul(thr, att);
                o = no;
                d = nd;
            } else {
                result = emitted(m);
                ended = true;
            }
        }
    } else {
        result = v_mul(thr, background(d));
        ended = true;
    }
    // Russian roulette / finish / regenerate, in the megakernel's order: RR is drawn before the next closest-hit
    // query whenever bounce >= min_bounces (CUDAKernels.h:113-121); a path that ends adds its radiance and the
    // pixel's next sample starts at once. The continuing lanes draw RR FIRST so that every way a path can end
    // (emission, sky, roulette, bounce limit) reaches the "path ended" block together: the block (accumulator
    // load, camera ray with its rejection loop) is issued once per warp, not once per reason. A fresh path starts
    // at bounce 0 < min_bounces (crt_render enforces min_bounces >= 1 and max_bounces >= 1), so it draws no RR here.
    if (!ended) {
        bounce++;
        if (bounce >= a.rc.max_bounces) {
            ended = true;                                             // exhausted: black
        } else if (bounce >= a.rc.min_bounces) {
            const float p = fminf(fmaxf(thr.x, fmaxf(thr.y, thr.z)), a.rc.rr_max_prob);
            if (rng_uniform(rng) > p) ended = true;
            else thr = v_scale(thr, f_rcp(p));
        }
    }
    keep = true;
    if (ended) {
        if (!touched) { acc = a.fb.accum[pix]; touched = true; }
        acc.x = f_add(acc.x, result.x); acc.y = f_add(acc.y, result.y); acc.z = f_add(acc.z, result.z);
        sample++;
        if (sample >= a.rc.spp) {
            keep = false;
        } else {
            const int x = (int)(pix % (uint32_t)a.rc.width), y = (int)(pix / (uint32_t)a.rc.width);
            camera_ray(a.cam, x, y, a.rc.width, a.rc.height, rng, o, d);
            thr = mk3(1.f, 1.f, 1.f);
            bounce = 0;
        }
    }
    return keep;
}

// ---- shade: consume the hit, continue / finish / regenerate, Russian roulette, compact -----------------
#ifndef CRT_SHADE_MIN_BLOCKS
#define CRT_SHADE_MIN_BLOCKS 4   // 64 registers, no spills: 6 % faster than 5 CTAs of 48 registers since the one-pass shade_step
#endif
__global__ void __launch_bounds__(256, CRT_SHADE_MIN_BLOCKS) k_wf_shade(RenderArgs a, WfView w, uint32_t* queue, uint32_t* qcount, int cur, int par) {
    const uint32_t count = qcount[cur];
    const uint32_t rounded = (count + 31u) & ~31u;
    const size_t n = (size_t)a.rc.width * a.rc.height;
    const uint32_t first = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    uint32_t pix_next = first < count ? queue[first] : QUEUE_DEAD;
    for (uint32_t i = first; i < rounded; i += stride) {
        bool keep = false;
        const uint32_t pix = pix_next;
        // the thread's next entry: its queue word is loaded now and its path state prefetched below, so that the
        // next trip does not start with a chain of DRAM latencies (queue -> state -> primitive record)
        pix_next = (i + stride < count) ? queue[i + stride] : QUEUE_DEAD;
        if (pix != QUEUE_DEAD) {
            const float4 A = CRT_LD_STATE_S(&w.A[pix]), B = CRT_LD_STATE_S(&w.B[pix]), C = CRT_LD_STATE_S(&w.C[pix]);
            if (C.x == PARKED_T) continue;   // a parked pixel waits for extend to finish its ray: nothing is touched
#if CRT_SHADE_PREFETCH
            asm volatile("prefetch.global.L2 [%0];" :: "l"(a.fb.accum + pix));   // read only when a path ends, late in the body
            if (pix_next != QUEUE_DEAD) {
                asm volatile("prefetch.global.L2 [%0];" :: "l"(w.A + pix_next));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(w.B + pix_next));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(w.C + pix_next));
#pragma unroll
                for (int k = 0; k < 6; k++) asm volatile("prefetch.global.L2 [%0];" :: "l"(a.fb.rng + (size_t)k * n + pix_next));
            }
#endif
#if CRT_SHADE_PREFETCH_PRIM
            float2 next_hit = make_float2(-1.f, 0.f);   // (t, record index) of the thread's next entry, written by the extend before this launch
            if (pix_next != QUEUE_DEAD) next_hit = __ldcg(reinterpret_cast<const float2*>(&w.C[pix_next]));
#endif
            vec3 o = mk3(A.x, A.y, A.z), d = mk3(B.x, B.y, B.z), thr = mk3(fabsf(A.w), B.w, C.z);   // |A.w|: PARKED_SIGN may be left on it
            const uint32_t book = __float_as_uint(C.w);
            int sample = (int)(book & 0xffffffu), bounce = (int)(book >> 24);
            Rng rng = wf_load_rng(a.fb.rng, n, pix);
            bool touched = false;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            keep = shade_step(a, pix, o, d, thr, sample, bounce, rng, C.x, __float_as_uint(C.y), acc, touched);
#if CRT_SHADE_PREFETCH_PRIM
            if (next_hit.x >= 0.f) {
                const char* rec = reinterpret_cast<const char*>(a.scene.prims + __float_as_uint(next_hit.y));
                asm volatile("prefetch.global.L1 [%0];" :: "l"(rec));
                asm volatile("prefetch.global.L1 [%0];" :: "l"(rec + 32));
            }
#endif
            if (touched) {
                if (!keep) {
                    acc.w += (float)a.rc.spp;
                    if (!(a.flags & CRT_RENDER_NO_TONEMAP))
                        a.fb.rgba8[pix] = pack_rgba8(v_scale(mk3(acc.x, acc.y, acc.z), f_div(1.f, acc.w)));
                }
                a.fb.accum[pix] = acc;
            }
            wf_store_rng(a.fb.rng, n, pix, rng);
            if (keep) {
#if CRT_STATE_STORE_CG
                __stcg(&w.A[pix], make_float4(o.x, o.y, o.z, thr.x));
                __stcg(&w.B[pix], make_float4(d.x, d.y, d.z, thr.y));
                __stcg(reinterpret_cast<float2*>(&w.C[pix]) + 1, make_float2(thr.z, __uint_as_float((uint32_t)sample | ((uint32_t)bounce << 24))));
#else
                w.A[pix] = make_float4(o.x, o.y, o.z, thr.x);
                w.B[pix] = make_float4(d.x, d.y, d.z, thr.y);
                float2* c2 = reinterpret_cast<float2*>(&w.C[pix]) + 1;
                *c2 = make_float2(thr.z, __uint_as_float((uint32_t)sample | ((uint32_t)bounce << 24)));
#endif
            } else {
                queue[i] = QUEUE_DEAD;   // all samples of this pixel are done
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        qcount[2] = 0u;                // fetch cursor of the next extend
        qcount[4 + (par ^ 1)] = 0u;    // the parked-ray list the extend before this shade has consumed: the next extend fills it
    }
}

// ---- shade, fed by the bulk-copy engine ---------------------------------------------------------------
// The queue stays in pixel order and (until pixels start to finish, ~85 % of a 2000-spp frame) without holes, so
// the 32 entries a warp takes per trip are 32 CONSECUTIVE pixels: their A, B, C records are three contiguous
// 512-byte runs and their six RNG words six contiguous 128-byte runs. One lane hands those nine runs to the
// bulk asynchronous copy engine (cp.async.bulk global -> shared, completion counted on an mbarrier; the 1-D form
// of TMA) ONE TRIP AHEAD, into the warp's 2304-byte stage: a trip first moves its entries from the stage to
// registers, then refills the stage for the next trip and computes while the copy runs. The per-thread version above issues 9 dependent loads + 9 L2 prefetches per entry and
// still starts every trip with an L2 round trip; here the state of the next trip is already on the SM when
// the trip begins, and the load/prefetch/address instructions are gone. Each warp owns its stages and
// barriers: no CTA-wide synchronisation, warps drift apart as their rejection loops dictate.
// A warp whose 32 entries are not one aligned run (holes after pixels finished, the ragged end of the queue)
// falls back to per-thread loads for that trip (with the L2 prefetch one trip ahead). Results leave through plain
// coalesced stores. Same arithmetic per pixel: the frame's bits do not change (the parity tests run both kernels).
#define CRT_SHADE_STAGE_BYTES 2304u   // 3 x 512 (A, B, C) + 6 x 128 (RNG planes)
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0u;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    // a bulk copy that never completes would hang the GPU: trap instead (a few seconds of polling)
    for (uint32_t spin = 0; !mbar_try_wait(bar, parity); spin++)
        if (spin > (1u << 26)) __trap();
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

__global__ void __launch_bounds__(256, CRT_SHADE_MIN_BLOCKS) k_wf_shade_bulk(RenderArgs a, WfView w, uint32_t* queue, uint32_t* qcount, int cur, int par) {
    __shared__ __align__(128) unsigned char stages[8][CRT_SHADE_STAGE_BYTES];   // one stage per warp
    __shared__ __align__(8) unsigned long long bars[8];
    const uint32_t count = qcount[cur];
    const uint32_t rounded = (count + 31u) & ~31u;
    const size_t n = (size_t)a.rc.width * a.rc.height;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t first = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    const bool planes_aligned = (n & 3u) == 0u;   // every RNG plane starts on a 16-byte boundary
    const uint32_t bar = smem_addr(&bars[warp]), dst = smem_addr(&stages[warp][0]);
    const unsigned char* sb = &stages[warp][0];
    if (lane == 0) {
        mbar_init(bar, 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    uint32_t phase = 0u;     // parity of the barrier phase the next bulk load completes (warp-uniform)
    auto queue_word = [&](unsigned long long i) -> uint32_t { return i < (unsigned long long)count ? queue[i] : QUEUE_DEAD; };
    // Start the copy of the 32 entries `qw` (one per lane) into the warp's stage when they are one aligned run of
    // pixels; otherwise prefetch them into L2 for the per-thread loads of the next trip. True = the stage is filling.
    auto issue = [&](uint32_t qw) -> bool {
        const uint32_t qbase = __shfl_sync(0xffffffffu, qw, 0);
        const bool run = __all_sync(0xffffffffu, qw != QUEUE_DEAD && qw == qbase + lane) && planes_aligned && (qbase & 3u) == 0u;
        if (run) {
            if (lane == 0) {
                mbar_expect_tx(bar, CRT_SHADE_STAGE_BYTES);
                bulk_g2s(dst, w.A + qbase, 512u, bar);
                bulk_g2s(dst + 512u, w.B + qbase, 512u, bar);
                bulk_g2s(dst + 1024u, w.C + qbase, 512u, bar);
#pragma unroll
                for (uint32_t k = 0; k < 6u; k++) bulk_g2s(dst + 1536u + 128u * k, a.fb.rng + (size_t)k * n + qbase, 128u, bar);
            }
        } else if (qw != QUEUE_DEAD) {
            asm volatile("prefetch.global.L2 [%0];" :: "l"(w.A + qw));
            asm volatile("prefetch.global.L2 [%0];" :: "l"(w.B + qw));
            asm volatile("prefetch.global.L2 [%0];" :: "l"(w.C + qw));
#pragma unroll
            for (int k = 0; k < 6; k++) asm volatile("prefetch.global.L2 [%0];" :: "l"(a.fb.rng + (size_t)k * n + qw));
        }
        if (qw != QUEUE_DEAD) asm volatile("prefetch.global.L2 [%0];" :: "l"(a.fb.accum + qw));   // read only when a path ends
        return run;
    };
    // the warp's queue words of this trip and the next (64-bit index: first + 2*stride may pass 2^32)
    uint32_t q_cur = queue_word(first), q_next = queue_word((unsigned long long)first + stride);
    bool staged = first < rounded ? issue(q_cur) : false;
    for (uint32_t i = first; i < rounded;) {
        const uint32_t pix = q_cur;
        const unsigned long long i_next = (unsigned long long)i + stride;
        const uint32_t q_after = queue_word(i_next + stride);
        float4 A, B, C;
        Rng rng;
        A = B = C = make_float4(0.f, 0.f, 0.f, 0.f);
        rng.v0 = rng.v1 = rng.v2 = rng.v3 = rng.v4 = rng.d = 0u;
        if (staged) {
            mbar_wait(bar, phase);
            phase ^= 1u;
            A = reinterpret_cast<const float4*>(sb)[lane];
            B = reinterpret_cast<const float4*>(sb + 512)[lane];
            C = reinterpret_cast<const float4*>(sb + 1024)[lane];
            const uint32_t* sr = reinterpret_cast<const uint32_t*>(sb + 1536);
            rng.v0 = sr[lane]; rng.v1 = sr[32 + lane]; rng.v2 = sr[64 + lane];
            rng.v3 = sr[96 + lane]; rng.v4 = sr[128 + lane]; rng.d = sr[160 + lane];
        } else if (pix != QUEUE_DEAD) {
            A = w.A[pix]; B = w.B[pix]; C = w.C[pix];
            rng = wf_load_rng(a.fb.rng, n, pix);
        }
        // every lane holds its entry in registers: the stage is free again, and the next trip's entries start to
        // arrive while this trip computes
        __syncwarp();
        staged = i_next < (unsigned long long)rounded ? issue(q_next) : false;
        if (pix != QUEUE_DEAD && C.x != PARKED_T) {   // a parked pixel waits for extend to finish its ray: nothing is touched
            vec3 o = mk3(A.x, A.y, A.z), d = mk3(B.x, B.y, B.z), thr = mk3(fabsf(A.w), B.w, C.z);   // |A.w|: PARKED_SIGN may be left on it
            const uint32_t book = __float_as_uint(C.w);
            int sample = (int)(book & 0xffffffu), bounce = (int)(book >> 24);
            bool touched = false;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            const bool keep = shade_step(a, pix, o, d, thr, sample, bounce, rng, C.x, __float_as_uint(C.y), acc, touched);
            if (touched) {
                if (!keep) {
                    acc.w += (float)a.rc.spp;
                    if (!(a.flags & CRT_RENDER_NO_TONEMAP))
                        a.fb.rgba8[pix] = pack_rgba8(v_scale(mk3(acc.x, acc.y, acc.z), f_div(1.f, acc.w)));
                }
                a.fb.accum[pix] = acc;
            }
            wf_store_rng(a.fb.rng, n, pix, rng);
            if (keep) {
                w.A[pix] = make_float4(o.x, o.y, o.z, thr.x);
                w.B[pix] = make_float4(d.x, d.y, d.z, thr.y);
                float2* c2 = reinterpret_cast<float2*>(&w.C[pix]) + 1;
                *c2 = make_float2(thr.z, __uint_as_float((uint32_t)sample | ((uint32_t)bounce << 24)));
            } else {
                queue[i] = QUEUE_DEAD;   // all samples of this pixel are done
            }
        }
        q_cur = q_next;
        q_next = q_after;
        if (i_next >= (unsigned long long)rounded) break;
        i = (uint32_t)i_next;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        qcount[2] = 0u;                // fetch cursor of the next extend
        qcount[4 + (par ^ 1)] = 0u;    // the parked-ray list the extend before this shade has consumed: the next extend fills it
    }
}

// ---- tail: the last pixels of a frame, a thread per pixel for many iterations at a time ------------------
// A pixel's samples are sequential (its cuRAND stream), so the frame ends with a few pixels (those with the
// longest paths) that still owe hundreds of iterations. With a short queue an iteration costs what its longest
// ray costs (~0.1 ms, latency bound) while tracing almost nothing: measured on the Cornell+blob frame, the
// iterations with fewer than 150 k live pixels per part were 4 % of the rays and 10 % of the frame time.
// Once a part's queue is that short the host stops iterating it: after one closing extend (nothing parked), this
// kernel takes every remaining pixel with one thread that alternates shade_step and a whole closest-hit query
// (the megakernel's loop, entered in the middle of a path) for up to `max_iters` path segments; pixels that are
// not finished by then are written back in the "after extend" form, the queue is compacted and the next stage
// runs on what is left. Only the first `lanes` lanes of a warp take a pixel: the fewer pixels are left, the more
// warps they are spread over, because the latency of an iteration in a warp whose lanes are all at different
// points of different paths grows with the number of lanes (their branches are issued one after the other).
// Same operations in the same order per pixel as the wavefront iterations, so the frame's bits do not change.
#ifndef CRT_TAIL_THREADS
#define CRT_TAIL_THREADS 128
#endif
// The tail keeps its whole traversal stack in local memory: measured 1.7 % faster per frame than the shared-memory form
// (-DCRT_TAIL_SMEM_STACK=4). That form is also the build in which the loop below stalled on the B200 before its
// rendezvous got an explicit __syncwarp() (ptxas had compiled the loop-top ballots at the back edge
// without a divergence check).
#ifndef CRT_TAIL_SMEM_STACK
#define CRT_TAIL_SMEM_STACK 0
#endif
__global__ void __launch_bounds__(CRT_TAIL_THREADS) k_wf_tail(RenderArgs a, WfView w, uint32_t* queue, const uint32_t* qcount, int cur, int lanes, int max_iters, int shade_lanes) {
#if CRT_TAIL_SMEM_STACK > 0
    __shared__ uint2 sstack[CRT_TAIL_SMEM_STACK][CRT_TAIL_THREADS];
    TravStackShared<CRT_TAIL_SMEM_STACK, CRT_TAIL_THREADS> st;
    st.slice = &sstack[0][threadIdx.x];
#else
    TravStack st;
#endif
    const uint32_t count = qcount[cur];
    const size_t n = (size_t)a.rc.width * a.rc.height;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, warps = (gridDim.x * blockDim.x) >> 5;
    uint32_t rays = 0;
    for (uint32_t base = warp * (uint32_t)lanes; base < count; base += warps * (uint32_t)lanes) {
        const uint32_t i = base + lane;
        uint32_t pix = QUEUE_DEAD;
        if (lane < (uint32_t)lanes && i < count) pix = queue[i];
        bool alive = pix != QUEUE_DEAD;
        vec3 thr = mk3(0.f, 0.f, 0.f);
        int sample = 0, bounce = 0, it = 0;
        Rng rng = {};
        float hit_t = -1.f;
        uint32_t hit_prim = 0u;
        bool touched = false, need_shade = true;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        Traverser tr;
        tr.o = mk3(0.f, 0.f, 0.f); tr.d = mk3(0.f, 0.f, 1.f);
        if (alive) {
            const float4 A = w.A[pix], B = w.B[pix], C = w.C[pix];
            tr.o = mk3(A.x, A.y, A.z); tr.d = mk3(B.x, B.y, B.z); thr = mk3(fabsf(A.w), B.w, C.z);
            const uint32_t book = __float_as_uint(C.w);
            sample = (int)(book & 0xffffffu); bounce = (int)(book >> 24);
            rng = wf_load_rng(a.fb.rng, n, pix);
            hit_t = C.x;
            hit_prim = __float_as_uint(C.y);
        }
        // The lanes of the warp advance their traversals together, one wide-node visit per trip (Traverser::step);
        // a lane whose query has finished runs its shade step at the top of the next trip and joins the others
        // again with its next ray, instead of all lanes waiting for the warp's longest ray before any of them shades.
        // The shade step waits until `shade_lanes` lanes need it (or no lane has a node left to visit), so that its
        // long, branchy body is issued for several lanes at a time (the dynamic-fetch trade of k_wf_extend).
#if CRT_DEBUG_WATCHDOG
        uint32_t wd_trips = 0;
#endif
        while (true) {
#if CRT_DEBUG_WATCHDOG
            if (++wd_trips > (1u << 22)) {
                printf("WATCHDOG k_wf_tail blk %d thr %d alive %d need_shade %d pix %u sample %d bounce %d it %d sp %d ng %08x %08x rec.t %g o %g %g %g d %g %g %g lanes %d count %u\n",
                       blockIdx.x, threadIdx.x, (int)alive, (int)need_shade, pix, sample, bounce, it, tr.sp, tr.ng_x, tr.ng_y, tr.rec.t, tr.o.x, tr.o.y, tr.o.z, tr.d.x, tr.d.y, tr.d.z, lanes, count);
                break;
            }
#endif
            __syncwarp();   // the two ballots below are the loop's rendezvous: make the reconvergence explicit (tail stall)
            const unsigned want_shade = __ballot_sync(0xffffffffu, alive && need_shade);
            const unsigned stepping = __ballot_sync(0xffffffffu, alive && !need_shade);
            if ((want_shade | stepping) == 0u) break;
            if (alive && need_shade && (__popc(want_shade) >= shade_lanes || stepping == 0u)) {
                vec3 o = tr.o, d = tr.d;
                if (shade_step(a, pix, o, d, thr, sample, bounce, rng, hit_t, hit_prim, acc, touched)) {
                    tr.begin(o, d, 0.001f, INFINITY);
                    trace_giants<false>(a.scene, o, d, 0.001f, tr.rec, nullptr);
                    rays++;
                    need_shade = false;
                } else {
                    acc.w += (float)a.rc.spp;
                    if (!(a.flags & CRT_RENDER_NO_TONEMAP))
                        a.fb.rgba8[pix] = pack_rgba8(v_scale(mk3(acc.x, acc.y, acc.z), f_div(1.f, acc.w)));
                    a.fb.accum[pix] = acc;
                    wf_store_rng(a.fb.rng, n, pix, rng);
                    queue[i] = QUEUE_DEAD;
                    alive = false;
                }
            }
            __syncwarp();
            if (alive && !need_shade && tr.step<false>(a.scene, st, nullptr)) {
                const bool hit = tr.rec.id >= 0;
                hit_t = hit ? tr.rec.t : -1.f;
                hit_prim = tr.rec.prim;
                need_shade = true;
                if (++it >= max_iters) {   // the state k_wf_extend leaves behind: ray, throughput, book-keeping and the hit just found
                    w.A[pix] = make_float4(tr.o.x, tr.o.y, tr.o.z, thr.x);
                    w.B[pix] = make_float4(tr.d.x, tr.d.y, tr.d.z, thr.y);
                    w.C[pix] = make_float4(hit_t, __uint_as_float(hit_prim), thr.z, __uint_as_float((uint32_t)sample | ((uint32_t)bounce << 24)));
                    if (touched) a.fb.accum[pix] = acc;
                    wf_store_rng(a.fb.rng, n, pix, rng);
                    alive = false;
                }
            }
        }
    }
    rays = __reduce_add_sync(0xffffffffu, rays);
    if ((threadIdx.x & 31) == 0 && rays) atomicAdd(&a.fb.counters[0], (unsigned long long)rays);
}

cudaError_t wavefront_create(WavefrontState** out, int width, int height) {
    WavefrontState* s = new WavefrontState();
    s->n = (size_t)width * height;
    if (const char* v = getenv("CRT_WF_PARTS")) s->parts = atoi(v);
    if (s->parts < 1) s->parts = 1;
    if (s->parts > CRT_WF_MAX_PARTS) s->parts = CRT_WF_MAX_PARTS;
    cudaError_t e = cudaMalloc(&s->A, s->n * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc(&s->B, s->n * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc(&s->C, s->n * sizeof(float4));
    const size_t tiles = (s->n + CRT_COMPACT_TILE - 1) / CRT_COMPACT_TILE;
    for (int p = 0; p < s->parts && e == cudaSuccess; p++) {
        WfPart& q = s->part[p];
        // part 0 can hold every pixel (a timed render runs as one part), the others their share of the tiles
        const size_t cap = p == 0 ? s->n : ((tiles + (size_t)s->parts - 1) / (size_t)s->parts) * CRT_COMPACT_TILE;
        e = cudaMalloc(&q.queue[0], cap * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaMalloc(&q.queue[1], cap * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaMalloc(&q.qcount, 8 * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaMalloc(&q.tile_count, (tiles + 1) * sizeof(uint32_t));
        if (e == cudaSuccess) e = cudaMallocHost(&q.h_qcount, 4 * sizeof(uint32_t));
        if (e == cudaSuccess && p > 0) e = cudaEventCreateWithFlags(&q.ev_done, cudaEventDisableTiming);
    }
    for (int p = 1; p < s->parts && e == cudaSuccess; p++) e = cudaStreamCreateWithFlags(&s->side_stream[p], cudaStreamNonBlocking);
    if (e == cudaSuccess && s->parts > 1) e = cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming);
    int be = 0, bs = 0;
    // CRT_EXTEND_CARVEOUT / CRT_SHADE_CARVEOUT / CRT_TAIL_CARVEOUT = percent of the SM's shared-memory capacity these kernels ask
    // the driver to set aside (cudaFuncAttributePreferredSharedMemoryCarveout): the rest of the 256 KB is L1, and extend
    // lives on its L1 hit rate. Extend needs 4 x 9 KB: 16 % (the 64 KB configuration) instead of the driver's default choice
    // (100 KB, by the occupancy ncu reports) measured +0.3 %, 100 % measured -2.6 %. -1 = leave the default.
    {
        const char* v = getenv("CRT_EXTEND_CARVEOUT");
        const int pct = v ? atoi(v) : 16;
        if (pct >= 0) {
        cudaFuncSetAttribute(k_wf_extend<false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        cudaFuncSetAttribute(k_wf_extend<true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        cudaFuncSetAttribute(k_wf_extend<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
        }
    }
    if (const char* v = getenv("CRT_SHADE_CARVEOUT")) cudaFuncSetAttribute(k_wf_shade, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(v));
    if (const char* v = getenv("CRT_TAIL_CARVEOUT")) cudaFuncSetAttribute(k_wf_tail, cudaFuncAttributePreferredSharedMemoryCarveout, atoi(v));
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&be, k_wf_extend<false, false>, CRT_EXTEND_THREADS, 0);
    if (const char* v = getenv("CRT_SHADE_BULK")) s->shade_bulk = atoi(v);
    if (const char* v = getenv("CRT_WF_GRAPH")) s->use_graph = atoi(v);
    if (const char* v = getenv("CRT_WF_BATCH")) s->batch = atoi(v) < 1 ? 1 : (atoi(v) > 16 ? 16 : atoi(v));
    if (const char* v = getenv("CRT_WF_BLOCKED")) s->blocked = atoi(v);
    if (const char* v = getenv("CRT_EXTEND_OCTANT")) s->extend_octant = atoi(v);
    if (e == cudaSuccess) e = s->shade_bulk ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bs, k_wf_shade_bulk, 256, 0)
                                            : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bs, k_wf_shade, 256, 0);
    if (const char* v = getenv("CRT_FETCH_LANES")) s->fetch_lanes = atoi(v);
    if (const char* v = getenv("CRT_EXTEND_REVERSE")) s->extend_reverse = atoi(v);
    if (const char* v = getenv("CRT_COOP")) s->coop = atoi(v);
    if (const char* v = getenv("CRT_SHADE_PERSISTENT")) s->shade_persistent = atoi(v);
    if (const char* v = getenv("CRT_POSTPONE_LANES")) s->postpone_lanes = atoi(v);
    if (const char* v = getenv("CRT_DRAIN_ROUNDS")) s->drain_rounds = atoi(v);
    if (const char* v = getenv("CRT_DRAIN_MIN_LIVE")) s->drain_min_live = atoll(v);
    if (const char* v = getenv("CRT_TAIL_LIVE")) s->tail_live = atoll(v);
    if (const char* v = getenv("CRT_TAIL_SHADE_LANES")) s->tail_shade_lanes = atoi(v);
    if (const char* v = getenv("CRT_TAIL_ITERS")) s->tail_iters = atoi(v) > 0 ? atoi(v) : 1;
    if (const char* v = getenv("CRT_TAIL_LANES_MIN")) s->tail_lanes_min = atoi(v) < 1 ? 1 : (atoi(v) > 32 ? 32 : atoi(v));
    { int bt = 0; if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bt, k_wf_tail, CRT_TAIL_THREADS, 0) == cudaSuccess && bt > 0) s->tail_blocks = bt; }
    if (s->fetch_lanes < 1) s->fetch_lanes = 1;
    if (s->fetch_lanes > 32) s->fetch_lanes = 32;
    if (const char* v = getenv("CRT_EXTEND_BLOCKS")) be = atoi(v);
    s->blocks_extend = be > 0 ? be : 4;
    s->blocks_shade = bs > 0 ? bs : 4;
    if (e != cudaSuccess) { wavefront_destroy(s); return e; }
    *out = s;
    return cudaSuccess;
}

void wavefront_destroy(WavefrontState* s) {
    if (!s) return;
    cudaFree(s->A); cudaFree(s->B); cudaFree(s->C);
    for (WfPart& q : s->part) {
        cudaFree(q.queue[0]); cudaFree(q.queue[1]); cudaFree(q.qcount); cudaFree(q.tile_count);
        cudaFree(q.suspended[0]); cudaFree(q.suspended[1]);
        cudaFree(q.oct_keys[0]); cudaFree(q.oct_keys[1]); cudaFree(q.oct_vals[0]); cudaFree(q.oct_vals[1]); cudaFree(q.oct_tmp);
        if (q.h_qcount) cudaFreeHost(q.h_qcount);
        if (q.ev_done) cudaEventDestroy(q.ev_done);
    }
    for (cudaStream_t st : s->side_stream) if (st) cudaStreamDestroy(st);
    if (s->ev_fork) cudaEventDestroy(s->ev_fork);
    for (cudaEvent_t ev : s->ev) if (ev) cudaEventDestroy(ev);
    for (WfGraph& g : s->graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
    delete s;
}

cudaError_t launch_render_wavefront(WavefrontState* ws, const RenderArgs& a, bool counted, int num_sms, cudaStream_t stream,
                                    uint32_t* launches, KernelTimes* times, bool* incomplete) {
    if (incomplete) *incomplete = false;
    if (!ws) return cudaErrorInvalidValue;
    WfView w;
    w.A = ws->A; w.B = ws->B; w.C = ws->C;
    const int grid_e = num_sms * ws->blocks_extend, grid_s = num_sms * ws->blocks_shade;
    const int batch = ws->batch;
    const bool timed = times != nullptr && (a.flags & CRT_RENDER_TIME_KERNELS);
    cudaError_t e = cudaSuccess;
    const int parts = ws->parts;
    const int ev_begin = 3 * 16 * CRT_WF_MAX_PARTS;
    if (timed && !ws->ev[0])
        for (int i = 0; i < ev_begin + 2 + 4 * CRT_WF_MAX_PARTS && e == cudaSuccess; i++) e = cudaEventCreate(&ws->ev[i]);
    if (e != cudaSuccess) return e;
    const size_t tiles_all = (ws->n + CRT_COMPACT_TILE - 1) / CRT_COMPACT_TILE;
    // rays are parked (see k_wf_extend) only by the plain kernel and only while the queue is longer than one
    // launch's worth of threads: a short queue runs out at the first fetch and nothing would be gained
    const bool suspend = ws->drain_rounds > 0 && !counted && !ws->coop;
    const long long min_live = ws->drain_min_live >= 0 ? ws->drain_min_live : (long long)grid_e * CRT_EXTEND_THREADS;
    BeginQueues bq = {};
    bq.parts = parts;
    for (int p = 0; p < parts; p++) {
        WfPart& q = ws->part[p];
        const size_t my_tiles = (tiles_all + (size_t)parts - 1 - (size_t)p) / (size_t)parts;
        size_t cap = my_tiles * CRT_COMPACT_TILE;
        if (my_tiles > 0 && (tiles_all - 1) % (size_t)parts == (size_t)p) cap -= tiles_all * CRT_COMPACT_TILE - ws->n;   // the partial last tile
        q.cap = (uint32_t)cap; q.live = q.cap; q.cur = 0; q.done = q.cap == 0u; q.tail = 0;
        for (int k = 0; k < 2 && e == cudaSuccess && suspend && !q.suspended[k]; k++)
            e = cudaMalloc(&q.suspended[k], (size_t)grid_e * CRT_EXTEND_THREADS * CRT_SUSPEND_WORDS * sizeof(uint2));
        // per-kernel times only mean something when the kernels run one after the other: a timed render keeps
        // the same launches (one per part) but puts them all on the caller's stream
        q.stream = (p == 0 || timed || ws->blocked) ? stream : ws->side_stream[p];
        bq.queue[p] = q.queue[0]; bq.qcount[p] = q.qcount; bq.cap[p] = q.cap;
    }
    if (e != cudaSuccess) return e;
    uint32_t nl = 0, compactions = 0, host_syncs = 0, nl_graph = 0;
    if (timed) cudaEventRecord(ws->ev[ev_begin], stream);
    k_wf_begin<<<grid_s, 256, 0, stream>>>(a, w, bq);
    if (timed) cudaEventRecord(ws->ev[ev_begin + 1], stream);
    nl++;
    e = cudaGetLastError();
    if (e == cudaSuccess && parts > 1 && !timed && !ws->blocked) {
        e = cudaEventRecord(ws->ev_fork, stream);
        for (int p = 1; p < parts && e == cudaSuccess; p++) e = cudaStreamWaitEvent(ws->side_stream[p], ws->ev_fork, 0);
    }
    // one iteration per path segment, plus the iterations in which a pixel's ray was parked (every parked ray
    // advances at least drain_rounds node visits per launch); a bound against a hang, not a tuning knob
    const long long max_iters = ((long long)a.rc.spp * a.rc.max_bounces + 1) * 256;
    long long it = 0;
    bool first = true;
    // CRT_WF_TRACE=1: one stderr line per batch (iteration, queue lengths, host wall clock) for tuning runs
    const bool trace = getenv("CRT_WF_TRACE") != nullptr;
    const bool debug_sync = getenv("CRT_WF_DEBUG_SYNC") != nullptr;   // blocking sync + one stderr line after every launch of the tail stages
    const auto trace_t0 = std::chrono::steady_clock::now();
    while (e == cudaSuccess && it < max_iters) {
        int active = 0;
        bool timed_part[CRT_WF_MAX_PARTS], tail_part[CRT_WF_MAX_PARTS];
        for (int p = 0; p < CRT_WF_MAX_PARTS; p++) { timed_part[p] = p < parts && !ws->part[p].done; tail_part[p] = p < parts && ws->part[p].tail != 0; }
        // iterations until the host looks at the queue lengths again: 16, but no more than can still be needed
        // when no ray gets parked (one iteration per path segment), and short batches for the stragglers after that
        const long long needed = (long long)a.rc.spp * a.rc.max_bounces + 1 - it;
        const int nb = needed >= batch ? batch : (needed > 0 ? (int)needed : 4);
        // ---- CUDA graph of the batch -------------------------------------------------------------------------
        // The launches of a batch depend on a handful of host-side facts (which queue is current, the parity of the
        // parked-ray lists, whether rays are parked at all) and on the kernel arguments. For ~85 % of a long frame
        // these repeat from batch to batch, so the batch is captured once (the second time its signature is seen:
        // a frame whose camera moves every frame never pays for an instantiation) and replayed with ONE launch:
        // 70 launches + 2 stream joins become one cudaGraphLaunch and one synchronisation.
        bool graphable = ws->use_graph && !timed && !counted && !ws->coop && !ws->extend_octant;
        for (int p = 0; p < parts; p++) if (!ws->part[p].done && ws->part[p].tail) graphable = false;   // tail grids follow the queue length
        WfGraph* graph = nullptr;
        bool capture = false;
        if (graphable) {
            std::vector<unsigned char> sig(sizeof(RenderArgs) + sizeof(WfView) + 64, 0);
            memcpy(sig.data(), &a, sizeof(RenderArgs));
            memcpy(sig.data() + sizeof(RenderArgs), &w, sizeof(WfView));
            unsigned char* t = sig.data() + sizeof(RenderArgs) + sizeof(WfView);
            *t++ = (unsigned char)nb; *t++ = (unsigned char)parts; *t++ = (unsigned char)ws->blocked; *t++ = (unsigned char)ws->shade_bulk;
            *t++ = (unsigned char)ws->fetch_lanes; *t++ = (unsigned char)ws->drain_rounds; *t++ = (unsigned char)ws->extend_reverse; *t++ = (unsigned char)ws->shade_persistent;
            for (int p = 0; p < parts; p++) {
                const WfPart& q = ws->part[p];
                *t++ = (unsigned char)q.done; *t++ = (unsigned char)q.cur; *t++ = (unsigned char)(q.extend_launches & 1u);
                *t++ = (unsigned char)((suspend && (long long)q.live >= min_live) ? 1 : 0);
            }
            memcpy(t, &stream, sizeof(stream));
            for (WfGraph& g : ws->graphs) if (g.sig == sig) { graph = &g; break; }
            if (!graph) {
                if (ws->graphs.size() >= 32) {   // bounded: drop the oldest
                    if (ws->graphs.front().exec) cudaGraphExecDestroy(ws->graphs.front().exec);
                    ws->graphs.erase(ws->graphs.begin());
                }
                ws->graphs.emplace_back();
                graph = &ws->graphs.back();
                graph->sig = sig;
            }
            graph->seen++;
            capture = !graph->exec && graph->seen >= 2;
        }
        const bool replay = graph && graph->exec;
        const bool emit = !replay;   // launches are issued (directly, or into the capture); bookkeeping always runs
        if (capture) {
            e = cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal);
            if (e == cudaSuccess && parts > 1 && !ws->blocked) {
                e = cudaEventRecord(ws->ev_fork, stream);
                for (int p = 1; p < parts && e == cudaSuccess; p++) e = cudaStreamWaitEvent(ws->side_stream[p], ws->ev_fork, 0);
            }
            if (e != cudaSuccess) break;
        }
        const int slots = nb * parts;
        for (int k = 0; k < slots; k++) {
            // iteration-major: the parts alternate (their kernels overlap on two streams); CRT_WF_BLOCKED: part-major
            const int b = ws->blocked ? k % nb : k / parts, p = ws->blocked ? k / nb : k % parts;
            {
                WfPart& q = ws->part[p];
                if (q.done) continue;
                cudaStream_t st = q.stream;
                if (q.tail) {   // one tail stage per batch (the first one after a closing extend that also completes the parked rays)
                    if (b > 0) continue;
                    if (timed) cudaEventRecord(ws->ev[ev_begin + 2 + 2 * p], st);
                    if (q.tail == 1) {
                        const int par = (int)(q.extend_launches++ & 1u);
                        k_wf_extend<false, false><<<grid_e, CRT_EXTEND_THREADS, 0, st>>>(a.scene, w, q.queue[q.cur], q.qcount, q.cur, a.fb.counters, ws->fetch_lanes, 0, q.suspended[par], q.suspended[par ^ 1], par, ws->extend_reverse);
                        q.tail = 2;
                        nl++;
                        if (debug_sync) { const cudaError_t de = cudaStreamSynchronize(st); fprintf(stderr, "wf_debug it=%lld part %d closing extend done (%s), live %u\n", it, p, cudaGetErrorName(de), q.live); }
                    }
                    // spread the pixels over the warps the machine holds (shared by the parts that are in their tail)
                    int tail_parts = 0;
                    for (int k2 = 0; k2 < parts; k2++) tail_parts += (!ws->part[k2].done && ws->part[k2].tail) ? 1 : 0;
                    const long long warps_resident = (long long)num_sms * ws->tail_blocks * (CRT_TAIL_THREADS / 32) / (tail_parts > 0 ? tail_parts : 1);
                    long long lanes = ((long long)q.live + warps_resident - 1) / (warps_resident > 0 ? warps_resident : 1);
                    if (lanes < ws->tail_lanes_min) lanes = ws->tail_lanes_min;
                    if (lanes > 32) lanes = 32;
                    const long long tail_warps = ((long long)q.live + lanes - 1) / lanes;
                    const int tail_grid = (int)((tail_warps * 32 + CRT_TAIL_THREADS - 1) / CRT_TAIL_THREADS);
                    const int shade_lanes = (int)(lanes * ws->tail_shade_lanes / 32) > 1 ? (int)(lanes * ws->tail_shade_lanes / 32) : 1;
                    k_wf_tail<<<tail_grid, CRT_TAIL_THREADS, 0, st>>>(a, w, q.queue[q.cur], q.qcount, q.cur, (int)lanes, ws->tail_iters, shade_lanes);
                    if (timed) cudaEventRecord(ws->ev[ev_begin + 3 + 2 * p], st);
                    nl++;
                    if (debug_sync) { const cudaError_t de = cudaStreamSynchronize(st); fprintf(stderr, "wf_debug it=%lld part %d tail stage done (%s), live %u lanes %d grid %d\n", it, p, cudaGetErrorName(de), q.live, (int)lanes, tail_grid); }
                    continue;
                }
                // shade: grid-stride over 148 x occupancy CTAs (CRT_SHADE_PERSISTENT=0: one CTA per 256 queue entries)
                const int shade_grid = ws->shade_persistent ? grid_s : (int)((q.live + 255u) / 256u);
                const int drain = (suspend && (long long)q.live >= min_live) ? ws->drain_rounds : 0;
                const int par = (int)(q.extend_launches++ & 1u);
                nl += 2;
                if (!emit) continue;
                uint32_t* equeue = q.queue[q.cur];
                if (ws->extend_octant && q.live > 0u) {   // untimed by the extend events: its cost is reported by the frame time only
                    size_t need = 0;
                    cub::DeviceRadixSort::SortPairs(nullptr, need, q.oct_keys[0], q.oct_keys[1], q.oct_vals[0], q.oct_vals[1], (int)q.live, 0, 4, st);
                    if (!q.oct_keys[0]) {
                        for (int k2 = 0; k2 < 2; k2++) { cudaMalloc(&q.oct_keys[k2], (size_t)q.cap * 4); cudaMalloc(&q.oct_vals[k2], (size_t)q.cap * 4); }
                    }
                    if (need > q.oct_tmp_bytes) { cudaFree(q.oct_tmp); cudaMalloc(&q.oct_tmp, need); q.oct_tmp_bytes = need; }
                    k_octant_keys<<<grid_s, 256, 0, st>>>(q.queue[q.cur], q.qcount, q.cur, w.B, q.oct_keys[0], q.oct_vals[0], q.live);
                    cub::DeviceRadixSort::SortPairs(q.oct_tmp, need, q.oct_keys[0], q.oct_keys[1], q.oct_vals[0], q.oct_vals[1], (int)q.live, 0, 4, st);
                    equeue = q.oct_vals[1];
                }
                if (timed) cudaEventRecord(ws->ev[3 * (b * parts + p)], st);
                if (counted) k_wf_extend<true, true><<<grid_e, CRT_EXTEND_THREADS, 0, st>>>(a.scene, w, equeue, q.qcount, q.cur, a.fb.counters, ws->fetch_lanes, drain, q.suspended[par], q.suspended[par ^ 1], par, ws->extend_reverse);
                else if (ws->coop) k_wf_extend<false, true><<<grid_e, CRT_EXTEND_THREADS, 0, st>>>(a.scene, w, equeue, q.qcount, q.cur, a.fb.counters, ws->fetch_lanes, drain, q.suspended[par], q.suspended[par ^ 1], par, ws->extend_reverse);
                else k_wf_extend<false, false><<<grid_e, CRT_EXTEND_THREADS, 0, st>>>(a.scene, w, equeue, q.qcount, q.cur, a.fb.counters, ws->fetch_lanes, drain, q.suspended[par], q.suspended[par ^ 1], par, ws->extend_reverse);
                if (timed) cudaEventRecord(ws->ev[3 * (b * parts + p) + 1], st);
                if (ws->shade_bulk) k_wf_shade_bulk<<<shade_grid, 256, 0, st>>>(a, w, q.queue[q.cur], q.qcount, q.cur, par);
                else k_wf_shade<<<shade_grid, 256, 0, st>>>(a, w, q.queue[q.cur], q.qcount, q.cur, par);
                if (timed) cudaEventRecord(ws->ev[3 * (b * parts + p) + 2], st);
            }
        }
        it += nb;
        // order-preserving compaction of the survivors into the other queue, then the new length comes back
        for (int p = 0; p < parts; p++) {
            WfPart& q = ws->part[p];
            if (q.done) continue;
            // a graph keeps its grids: sized for the part's capacity (CTAs past the queue's end return at once)
            const int tiles = (int)(((graphable ? q.cap : q.live) + CRT_COMPACT_TILE - 1) / CRT_COMPACT_TILE);
            nl += 3;
            compactions++;
            if (emit) {
                if (timed) cudaEventRecord(ws->ev[ev_begin + 2 + 2 * CRT_WF_MAX_PARTS + 2 * p], q.stream);
                k_compact_count<<<tiles, 256, 0, q.stream>>>(q.queue[q.cur], q.qcount, q.cur, q.tile_count);
                k_compact_scan<<<1, 1024, 0, q.stream>>>(q.tile_count, tiles, q.qcount, q.cur);
                k_compact_scatter<<<tiles, 256, 0, q.stream>>>(q.queue[q.cur], q.queue[q.cur ^ 1], q.qcount, q.cur, q.tile_count);
                if (timed) cudaEventRecord(ws->ev[ev_begin + 3 + 2 * CRT_WF_MAX_PARTS + 2 * p], q.stream);
            }
            q.cur ^= 1;
            if (emit) cudaMemcpyAsync(q.h_qcount, q.qcount, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, q.stream);
        }
        if (capture) {
            // join the side streams, close the capture, instantiate
            for (int p = 1; p < parts && e == cudaSuccess; p++) {
                WfPart& q = ws->part[p];
                if (q.stream == stream) continue;
                e = cudaEventRecord(q.ev_done, q.stream);
                if (e == cudaSuccess) e = cudaStreamWaitEvent(stream, q.ev_done, 0);
            }
            cudaGraph_t g = nullptr;
            const cudaError_t e2 = cudaStreamEndCapture(stream, &g);
            if (e == cudaSuccess) e = e2;
            if (e == cudaSuccess) e = cudaGraphInstantiate(&graph->exec, g, 0);
            if (g) cudaGraphDestroy(g);
            if (e != cudaSuccess) break;
        }
        if (graph && graph->exec) {
            e = cudaGraphLaunch(graph->exec, stream);
            nl_graph++;
            if (e == cudaSuccess) { e = cudaStreamSynchronize(stream); host_syncs++; }   // the side branches joined `stream` inside the graph
        }
        if (e == cudaSuccess) e = cudaGetLastError();
        for (int p = 0; p < parts && e == cudaSuccess; p++) {
            WfPart& q = ws->part[p];
            if (q.done) continue;
            if (!(graph && graph->exec)) {
                e = cudaStreamSynchronize(q.stream);
                if (p == 0 || q.stream != stream) host_syncs++;
                if (e != cudaSuccess) break;
            }
            q.live = q.h_qcount[q.cur];
            if (q.live == 0u) q.done = true; else active++;
        }
        if (e == cudaSuccess && timed) {
            float ms = 0.f;
            if (first) { cudaEventElapsedTime(&ms, ws->ev[ev_begin], ws->ev[ev_begin + 1]); times->begin_ms += ms; first = false; }
            for (int p = 0; p < parts; p++) {
                if (!timed_part[p]) continue;
                cudaEventElapsedTime(&ms, ws->ev[ev_begin + 2 + 2 * CRT_WF_MAX_PARTS + 2 * p], ws->ev[ev_begin + 3 + 2 * CRT_WF_MAX_PARTS + 2 * p]);
                times->compact_ms += ms;
                if (tail_part[p]) {
                    cudaEventElapsedTime(&ms, ws->ev[ev_begin + 2 + 2 * p], ws->ev[ev_begin + 3 + 2 * p]); times->tail_ms += ms;
                    continue;
                }
                for (int b = 0; b < nb; b++) {
                    const int k = 3 * (b * parts + p);
                    cudaEventElapsedTime(&ms, ws->ev[k], ws->ev[k + 1]); times->extend_ms += ms;
                    cudaEventElapsedTime(&ms, ws->ev[k + 1], ws->ev[k + 2]); times->shade_ms += ms;
                }
                times->extend_launches += nb;
                times->shade_launches += nb;
            }
        }
        // a part whose queue has become short leaves the wavefront iterations for the tail stages
        for (int p = 0; p < parts && !counted && !ws->coop; p++) {
            WfPart& q = ws->part[p];
            if (!q.done && !q.tail && (long long)q.live <= ws->tail_live) q.tail = 1;
        }
        if (trace)
            fprintf(stderr, "wf_trace it=%lld live=%u,%u,%u,%u ms=%.3f extend_ms=%.3f shade_ms=%.3f\n", it, ws->part[0].live, ws->part[1].live,
                    ws->part[2].live, ws->part[3].live, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - trace_t0).count(),
                    timed ? times->extend_ms : 0.0, timed ? times->shade_ms : 0.0);
        if (active == 0) break;
    }
    // every part's stream was synchronised by the host above, so work queued on `stream` after this call is ordered
    // behind the side stream's kernels as well. After an error the side streams may still hold work of the batch
    // that failed: wait for them too, so that nothing of this frame runs behind the caller's back.
    if (e != cudaSuccess)
        for (int p = 0; p < parts; p++) if (ws->part[p].stream) (void)cudaStreamSynchronize(ws->part[p].stream);
    if (e == cudaSuccess && incomplete) {   // the iteration bound was reached with pixels still live: the frame is NOT finished
        for (int p = 0; p < parts; p++) if (!ws->part[p].done) *incomplete = true;
    }
    if (launches) *launches += nl;
    if (times) { times->host_syncs += host_syncs; times->compactions += compactions; times->iterations += (uint32_t)it; times->graph_launches += nl_graph; }
    return e;
}

}  // namespace crt
