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
#include <cstdlib>

#include "crt_internal.h"
#include "crt_shade.h"

namespace crt {

#ifndef CRT_FETCH_LANES
#define CRT_FETCH_LANES 8
#endif

struct WavefrontState {
    float4* A = nullptr;
    float4* B = nullptr;
    float4* C = nullptr;
    uint32_t* queue[2] = {nullptr, nullptr};
    uint32_t* qcount = nullptr;     // [0],[1] queue sizes, [2] iteration scratch
    uint32_t* h_qcount = nullptr;   // pinned
    size_t n = 0;
    int blocks_extend = 0, blocks_shade = 0;
    int fetch_lanes = CRT_FETCH_LANES, postpone_lanes = CRT_POSTPONE_LANES;   // tuning knobs (env CRT_FETCH_LANES / CRT_POSTPONE_LANES)
    cudaEvent_t ev[3 * 16 + 2] = {};   // CRT_RENDER_TIME_KERNELS: begin pair + 3 events per iteration of a batch
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

// Warp-aggregated append: one atomic per warp, lanes ordered by ballot rank.
__device__ __forceinline__ void queue_push(uint32_t* queue, uint32_t* count, bool keep, uint32_t value) {
    const unsigned mask = __ballot_sync(0xffffffffu, keep);
    if (mask == 0u) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(count, (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (keep) queue[base + __popc(mask & ((1u << lane) - 1u))] = value;
}

// ---- begin: first camera ray of every pixel, queue = all pixels ---------------------------------
__global__ void __launch_bounds__(256) k_wf_begin(RenderArgs a, WfView w, uint32_t* queue, uint32_t* qcount) {
    const size_t n = (size_t)a.rc.width * a.rc.height;
    const size_t rounded = (n + 31) & ~(size_t)31;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < rounded; i += (size_t)gridDim.x * blockDim.x) {
        const bool valid = i < n;
        if (valid) {
            const int x = (int)(i % (size_t)a.rc.width), y = (int)(i / (size_t)a.rc.width);
            Rng rng = wf_load_rng(a.fb.rng, n, i);
            vec3 o, d;
            camera_ray(a.cam, x, y, a.rc.width, a.rc.height, rng, o, d);
            wf_store_rng(a.fb.rng, n, i, rng);
            w.A[i] = make_float4(o.x, o.y, o.z, 1.f);
            w.B[i] = make_float4(d.x, d.y, d.z, 1.f);
            w.C[i] = make_float4(0.f, 0.f, 1.f, __uint_as_float(0u));
            if (!(a.flags & CRT_RENDER_ACCUMULATE)) a.fb.accum[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            queue[i] = (uint32_t)i;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { qcount[0] = (uint32_t)n; qcount[1] = 0u; qcount[2] = 0u; }
}

// ---- extend: closest hit for every queued ray -------------------------------------------------------
// Persistent warps with dynamic fetch: a warp keeps up to 32 rays in flight and all its lanes advance
// their traversals one wide node at a time; when CRT_FETCH_LANES lanes have finished, they take the next
// rays from the queue together (one atomic per refill), so a long ray (one that wanders through the mesh)
// does not hold 31 finished lanes idle. qcount[2] is the shared fetch cursor (reset by begin / shade).
template <bool COUNT>
__global__ void __launch_bounds__(256) k_wf_extend(SceneView scene, WfView w, const uint32_t* queue, uint32_t* qcount, int cur,
                                                   unsigned long long* counters, int fetch_lanes, int postpone_lanes) {
    const uint32_t count = qcount[cur];
    const int lane = threadIdx.x & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    TraceCounters ctr;
    ctr.nodes = 0; ctr.prims = 0;
    Traverser tr;
    TravStack st;
    bool has_ray = false, exhausted = false;
    uint32_t pix = 0;
    while (true) {
        unsigned idle = __ballot_sync(0xffffffffu, !has_ray);
        if (!exhausted && __popc(idle) >= fetch_lanes) {
            const int n = __popc(idle), leader = __ffs(idle) - 1;
            uint32_t base = 0;
            if (lane == leader) base = atomicAdd(&qcount[2], (uint32_t)n);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (base + (uint32_t)n >= count) exhausted = true;
            const uint32_t i = base + (uint32_t)__popc(idle & lt_mask);
            if (!has_ray && i < count) {
                pix = queue[i];
                const float4 A = w.A[pix], B = w.B[pix];
                tr.begin(mk3(A.x, A.y, A.z), mk3(B.x, B.y, B.z), 0.001f, INFINITY);
                has_ray = true;
            }
            idle = __ballot_sync(0xffffffffu, !has_ray);
        }
        if (idle == 0xffffffffu) break;
        if (has_ray) {
            if (tr.step<COUNT>(scene, st, &ctr, postpone_lanes)) {
                float2* c = reinterpret_cast<float2*>(&w.C[pix]);
                *c = make_float2(tr.rec.id >= 0 ? tr.rec.t : -1.f, __uint_as_float(tr.rec.prim));
                has_ray = false;
            }
        }
    }
    if (COUNT) {
        unsigned long long nodes = ctr.nodes, prims = ctr.prims;
        for (int o = 16; o > 0; o >>= 1) { nodes += __shfl_xor_sync(0xffffffffu, nodes, o); prims += __shfl_xor_sync(0xffffffffu, prims, o); }
        if (lane == 0 && (nodes | prims)) { atomicAdd(&counters[2], nodes); atomicAdd(&counters[3], prims); }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        counters[0] += count;        // one ray per queued path; only this thread writes counters[0]
        qcount[cur ^ 1] = 0u;         // the other queue is free again: shade will append to it
    }
}

// ---- shade: consume the hit, continue / finish / regenerate, Russian roulette, compact -----------------
__global__ void __launch_bounds__(256) k_wf_shade(RenderArgs a, WfView w, const uint32_t* queue_in, uint32_t* queue_out,
                                                  uint32_t* qcount, int cur) {
    const uint32_t count = qcount[cur];
    const uint32_t rounded = (count + 31u) & ~31u;
    const size_t n = (size_t)a.rc.width * a.rc.height;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < rounded; i += gridDim.x * blockDim.x) {
        bool keep = false;
        uint32_t pix = 0;
        if (i < count) {
            pix = queue_in[i];
            const float4 A = w.A[pix], B = w.B[pix], C = w.C[pix];
            vec3 o = mk3(A.x, A.y, A.z), d = mk3(B.x, B.y, B.z), thr = mk3(A.w, B.w, C.z);
            const uint32_t book = __float_as_uint(C.w);
            int sample = (int)(book & 0xffffffu), bounce = (int)(book >> 24);
            Rng rng = wf_load_rng(a.fb.rng, n, pix);
            bool ended = false;
            vec3 result = mk3(0.f, 0.f, 0.f);
            if (!(C.x < 0.f)) {                                  // extend stores t = -1 for a miss
                HitRec h;
                h.t = C.x; h.prim = __float_as_uint(C.y); h.id = 0;
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
            if (!ended) bounce++;
            // Finish / regenerate / Russian roulette, in the megakernel's order: a path that ends
            // adds its radiance and the pixel's next sample starts at once; RR is drawn before the
            // next closest-hit query whenever bounce >= min_bounces (CUDAKernels.h:113-121).
            keep = true;
            bool touched = false;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            while (true) {
                if (ended) {
                    if (!touched) { acc = a.fb.accum[pix]; touched = true; }
                    acc.x = f_add(acc.x, result.x); acc.y = f_add(acc.y, result.y); acc.z = f_add(acc.z, result.z);
                    sample++;
                    if (sample >= a.rc.spp) { keep = false; break; }
                    const int x = (int)(pix % (uint32_t)a.rc.width), y = (int)(pix / (uint32_t)a.rc.width);
                    camera_ray(a.cam, x, y, a.rc.width, a.rc.height, rng, o, d);
                    thr = mk3(1.f, 1.f, 1.f);
                    bounce = 0;
                    result = mk3(0.f, 0.f, 0.f);
                    ended = false;
                }
                if (bounce >= a.rc.max_bounces) { ended = true; continue; }   // exhausted: black
                if (bounce >= a.rc.min_bounces) {
                    const float p = fminf(fmaxf(thr.x, fmaxf(thr.y, thr.z)), a.rc.rr_max_prob);
                    if (rng_uniform(rng) > p) { ended = true; continue; }
                    thr = v_scale(thr, f_rcp(p));
                }
                break;
            }
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
            }
        }
        queue_push(queue_out, &qcount[cur ^ 1], keep, pix);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) qcount[2] = 0u;   // fetch cursor of the next extend
}

cudaError_t wavefront_create(WavefrontState** out, int width, int height) {
    WavefrontState* s = new WavefrontState();
    s->n = (size_t)width * height;
    cudaError_t e = cudaMalloc(&s->A, s->n * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc(&s->B, s->n * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc(&s->C, s->n * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc(&s->queue[0], s->n * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&s->queue[1], s->n * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMalloc(&s->qcount, 4 * sizeof(uint32_t));
    if (e == cudaSuccess) e = cudaMallocHost(&s->h_qcount, 4 * sizeof(uint32_t));
    int be = 0, bs = 0;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&be, k_wf_extend<false>, 256, 0);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bs, k_wf_shade, 256, 0);
    if (const char* v = getenv("CRT_FETCH_LANES")) s->fetch_lanes = atoi(v);
    if (const char* v = getenv("CRT_POSTPONE_LANES")) s->postpone_lanes = atoi(v);
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
    cudaFree(s->A); cudaFree(s->B); cudaFree(s->C); cudaFree(s->queue[0]); cudaFree(s->queue[1]); cudaFree(s->qcount);
    if (s->h_qcount) cudaFreeHost(s->h_qcount);
    for (cudaEvent_t ev : s->ev) if (ev) cudaEventDestroy(ev);
    delete s;
}

cudaError_t launch_render_wavefront(WavefrontState* ws, const RenderArgs& a, bool counted, int num_sms, cudaStream_t stream,
                                    uint32_t* launches, KernelTimes* times) {
    if (!ws) return cudaErrorInvalidValue;
    WfView w;
    w.A = ws->A; w.B = ws->B; w.C = ws->C;
    const int grid_e = num_sms * ws->blocks_extend, grid_s = num_sms * ws->blocks_shade;
    const int batch = 16;
    const bool timed = times != nullptr && (a.flags & CRT_RENDER_TIME_KERNELS);
    cudaError_t e = cudaSuccess;
    if (timed && !ws->ev[0])
        for (int i = 0; i < 3 * batch + 2 && e == cudaSuccess; i++) e = cudaEventCreate(&ws->ev[i]);
    if (e != cudaSuccess) return e;
    uint32_t nl = 0;
    if (timed) cudaEventRecord(ws->ev[3 * batch], stream);
    k_wf_begin<<<grid_s, 256, 0, stream>>>(a, w, ws->queue[0], ws->qcount);
    if (timed) cudaEventRecord(ws->ev[3 * batch + 1], stream);
    nl++;
    int cur = 0;
    const long long max_iters = (long long)a.rc.spp * a.rc.max_bounces + 1;
    long long it = 0;
    bool first = true;
    e = cudaGetLastError();
    while (e == cudaSuccess && it < max_iters) {
        for (int b = 0; b < batch; b++) {
            if (timed) cudaEventRecord(ws->ev[3 * b], stream);
            if (counted) k_wf_extend<true><<<grid_e, 256, 0, stream>>>(a.scene, w, ws->queue[cur], ws->qcount, cur, a.fb.counters, ws->fetch_lanes, ws->postpone_lanes);
            else k_wf_extend<false><<<grid_e, 256, 0, stream>>>(a.scene, w, ws->queue[cur], ws->qcount, cur, a.fb.counters, ws->fetch_lanes, ws->postpone_lanes);
            if (timed) cudaEventRecord(ws->ev[3 * b + 1], stream);
            k_wf_shade<<<grid_s, 256, 0, stream>>>(a, w, ws->queue[cur], ws->queue[cur ^ 1], ws->qcount, cur);
            if (timed) cudaEventRecord(ws->ev[3 * b + 2], stream);
            nl += 2;
            cur ^= 1;
            it++;
        }
        e = cudaGetLastError();
        if (e != cudaSuccess) break;
        e = cudaMemcpyAsync(ws->h_qcount, ws->qcount, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
        if (e == cudaSuccess && timed) {
            float ms = 0.f;
            if (first) { cudaEventElapsedTime(&ms, ws->ev[3 * batch], ws->ev[3 * batch + 1]); times->shade_ms += ms; times->shade_launches++; first = false; }
            for (int b = 0; b < batch; b++) {
                cudaEventElapsedTime(&ms, ws->ev[3 * b], ws->ev[3 * b + 1]); times->extend_ms += ms;
                cudaEventElapsedTime(&ms, ws->ev[3 * b + 1], ws->ev[3 * b + 2]); times->shade_ms += ms;
            }
            times->extend_launches += batch;
            times->shade_launches += batch;
        }
        if (e == cudaSuccess && ws->h_qcount[cur] == 0u) break;
    }
    if (launches) *launches += nl;
    return e;
}

}  // namespace crt
