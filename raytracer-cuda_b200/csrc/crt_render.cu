// crt_render.cu — RNG init (block-cooperative XORWOW skip-ahead), megakernel path tracer, primary-id pass, tonemap (sm_100a).
//
// Replaces CUDAKernels::initRandState and CUDAKernels::render
// (reference: CudaRayTracer/src/CUDAKernels.h:18-26,147-166).
#include <cuda_runtime.h>
#include <curand_kernel.h>

#include <cstdlib>
#include <mutex>

#include "../../include/crt_api.h"

#include "crt_internal.h"
#include "crt_shade.h"

namespace crt {

// Pixel owned by a thread: 16x16 pixel tile per 256-thread block, each warp an 8x4 sub-tile so the
// 32 camera rays of a warp stay close together.
__device__ __forceinline__ bool tile_pixel(int width, int height, int& x, int& y) {
    const int tiles_x = (width + 15) >> 4;
    const int bx = blockIdx.x % tiles_x, by = blockIdx.x / tiles_x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    x = bx * 16 + (warp & 1) * 8 + (lane & 7);
    y = by * 16 + (warp >> 1) * 4 + (lane >> 3);
    return x < width && y < height;
}

// ---- RNG init: state[pixel] = curand_init(seed, subseq_base + pixel, 0) -----------------------
// cuRAND's own device routine does the 2^67-per-subsequence skip-ahead; only v[0..4] and d are
// kept (curand_uniform touches nothing else), as 6 coalesced planes instead of 48-byte AoS states.
__global__ void __launch_bounds__(256) k_rng_init(uint32_t* planes, int width, int height, unsigned long long seed,
                                                  unsigned long long base) {
    const size_t n = (size_t)width * height;
    for (size_t pix = (size_t)blockIdx.x * blockDim.x + threadIdx.x; pix < n; pix += (size_t)gridDim.x * blockDim.x) {
        curandStateXORWOW_t st;
        curand_init(seed, base + pix, 0, &st);
        planes[pix] = st.v[0];
        planes[n + pix] = st.v[1];
        planes[2 * n + pix] = st.v[2];
        planes[3 * n + pix] = st.v[3];
        planes[4 * n + pix] = st.v[4];
        planes[5 * n + pix] = st.d;
    }
}

// ---- the same states, computed cooperatively ------------------------------------------------------------------
// curand_init(seed, n, 0) is v(n) = P(n) v0: v0 depends on the seed only, and P(n) = product over the base-4 digits
// d_k of n of (A^(4^k))^(d_k), A = the XORWOW step matrix to the power 2^67 (one subsequence). cuRAND ships the
// 160 x 160 bit matrices A^(4^k) (precalc_xorwow_matrix[k], curand_precalc.h) and curand_init applies them per
// THREAD: ~16 matrix-vector products of 160 rows each for a 22-bit pixel index, 35 k instructions and a local-memory
// scratch vector per pixel (8.4 ms for a 2560x1440 frame, ncu: 4.06 G warp instructions, lg-throttle bound).
// All these matrices are powers of one matrix, so they commute and the 256 consecutive subsequences of a CTA are
//   v(n0 + r) = A^r v(n0),  r = sum e_k 4^k:   S[r + e 4^k] = A^(4^k) S[r + (e-1) 4^k]   for e = 1..3, k = 0..3
// i.e. 255 products per CTA after ONE generic v(n0), and each product is done by a WARP: lane l takes row
// 32 i + l of every state word i (5 rows per lane), then an xor butterfly over the lanes. 127 warp instructions per
// pixel instead of 1 100. Same bits as curand_init (tests: every pixel of full-size frames against k_rng_init, and
// the oracle's host restatement). d (the Weyl counter) does not move: a subsequence is a multiple of 2^32 steps.
__device__ __forceinline__ void xorwow_matvec_warp(const unsigned int* __restrict__ m, const uint32_t (&in)[5], uint32_t (&out)[5], int lane) {
    uint32_t a0 = 0u, a1 = 0u, a2 = 0u, a3 = 0u, a4 = 0u;
#pragma unroll
    for (int i = 0; i < 5; i++) {
        // row i*32 + lane of the matrix: what bit `lane` of state word i contributes (__curand_matvec). Branch-free (the row
        // is masked, not skipped), so the lanes reach the butterfly below without ever having diverged
        const uint32_t mask = 0u - ((in[i] >> lane) & 1u);
        const unsigned int* row = m + 5 * (i * 32 + lane);
        a0 ^= row[0] & mask; a1 ^= row[1] & mask; a2 ^= row[2] & mask; a3 ^= row[3] & mask; a4 ^= row[4] & mask;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a0 ^= __shfl_xor_sync(0xffffffffu, a0, o);
        a1 ^= __shfl_xor_sync(0xffffffffu, a1, o);
        a2 ^= __shfl_xor_sync(0xffffffffu, a2, o);
        a3 ^= __shfl_xor_sync(0xffffffffu, a3, o);
        a4 ^= __shfl_xor_sync(0xffffffffu, a4, o);
    }
    out[0] = a0; out[1] = a1; out[2] = a2; out[3] = a3; out[4] = a4;
}

__global__ void __launch_bounds__(256) k_rng_init_coop(uint32_t* planes, unsigned long long n_pixels, unsigned long long seed,
                                                       unsigned long long base) {
    __shared__ uint32_t S[256][5];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned long long first = (unsigned long long)blockIdx.x * 256ull;
    // seed scramble of _curand_init_inplace (curand_kernel.h)
    const unsigned int s0 = ((unsigned int)seed) ^ 0xaad26b49u;
    const unsigned int s1 = (unsigned int)(seed >> 32) ^ 0xf7dcefddu;
    const unsigned int t0 = 1099087573u * s0;
    const unsigned int t1 = 2591861531u * s1;
    const uint32_t d0 = 6615241u + t1 + t0;
    if (warp == 0) {   // v(n0): the generic digit loop of _skipahead_sequence, each product spread over the warp
        uint32_t v[5] = {123456789u + t0, 362436069u ^ t0, 521288629u + t1, 88675123u ^ t1, 5783321u + t0};
        unsigned long long p = base + first;
        int k = 0;
        while (p) {   // k stays below PRECALC_NUM_MATRICES - 1: the launcher sends larger subsequences to k_rng_init
            const unsigned int digit = (unsigned int)(p & PRECALC_BLOCK_MASK);
            for (unsigned int t = 0; t < digit; t++) xorwow_matvec_warp(precalc_xorwow_matrix[k], v, v, lane);
            p >>= PRECALC_BLOCK_SIZE;
            k++;
        }
        if (lane == 0) { S[0][0] = v[0]; S[0][1] = v[1]; S[0][2] = v[2]; S[0][3] = v[3]; S[0][4] = v[4]; }
    }
    __syncthreads();
#pragma unroll 1
    for (int k = 0; k < 4; k++) {
        const int span = 1 << (2 * k);   // S[0 .. span) are known
        for (int e = 1; e <= 3; e++) {
            for (int r = warp; r < span; r += 8) {
                const uint32_t* src = S[r + (e - 1) * span];
                const uint32_t in[5] = {src[0], src[1], src[2], src[3], src[4]};
                uint32_t out[5];
                xorwow_matvec_warp(precalc_xorwow_matrix[k], in, out, lane);
                if (lane == 0) { uint32_t* dst = S[r + e * span]; dst[0] = out[0]; dst[1] = out[1]; dst[2] = out[2]; dst[3] = out[3]; dst[4] = out[4]; }
            }
            __syncthreads();
        }
    }
    const unsigned long long pix = first + threadIdx.x;
    if (pix < n_pixels) {
        planes[pix] = S[threadIdx.x][0];
        planes[n_pixels + pix] = S[threadIdx.x][1];
        planes[2 * n_pixels + pix] = S[threadIdx.x][2];
        planes[3 * n_pixels + pix] = S[threadIdx.x][3];
        planes[4 * n_pixels + pix] = S[threadIdx.x][4];
        planes[5 * n_pixels + pix] = d0;
    }
}

cudaError_t launch_rng_init(uint32_t* rng_planes, int width, int height, unsigned long long seed,
                            unsigned long long subseq_base, int num_sms, cudaStream_t stream) {
    const size_t n = (size_t)width * height;
    // subsequences from 4^31 on leave cuRAND's precalculated matrices (its own routine then squares matrices per thread):
    // those, and CRT_RNG_INIT_CURAND=1 (the reference path of the tests), go through cuRAND's curand_init
    const bool in_tables = subseq_base < (1ull << 61) && (unsigned long long)n < (1ull << 61);
    if (in_tables && !getenv("CRT_RNG_INIT_CURAND")) {
        k_rng_init_coop<<<(unsigned int)((n + 255) / 256), 256, 0, stream>>>(rng_planes, (unsigned long long)n, seed, subseq_base);
        return cudaGetLastError();
    }
    int blocks = (int)((n + 255) / 256);
    const int cap = num_sms * 8;
    if (blocks > cap) blocks = cap;
    k_rng_init<<<blocks, 256, 0, stream>>>(rng_planes, width, height, seed, subseq_base);
    return cudaGetLastError();
}

__device__ __forceinline__ Rng load_rng(const uint32_t* planes, size_t n, size_t pix) {
    Rng r;
    r.v0 = planes[pix]; r.v1 = planes[n + pix]; r.v2 = planes[2 * n + pix];
    r.v3 = planes[3 * n + pix]; r.v4 = planes[4 * n + pix]; r.d = planes[5 * n + pix];
    return r;
}
__device__ __forceinline__ void store_rng(uint32_t* planes, size_t n, size_t pix, const Rng& r) {
    planes[pix] = r.v0; planes[n + pix] = r.v1; planes[2 * n + pix] = r.v2;
    planes[3 * n + pix] = r.v3; planes[4 * n + pix] = r.v4; planes[5 * n + pix] = r.d;
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- megakernel: one thread per pixel, all samples of the pixel in sequence --------------------
// Same per-pixel work as CUDAKernels::render, restructured as a single loop with one traversal
// call site: a lane whose path ended starts its next sample immediately, so every lane of a warp
// has a ray in flight until its pixel is finished.
// Traversal stack: all in local memory. 4 entries in shared memory (-DCRT_MEGA_SMEM_STACK=4, the layout that pays in
// k_wf_extend) measured slower on every config here: C3 161.0 vs 156.9 ms, C1 0.752 vs 0.741 ms, C5 2.334 vs 2.229 ms,
// terrain (16 spp) 124.9 vs 119.3 ms (scripts/r2_call27.sh): the megakernel's warps are register-, not L1-bound.
#ifndef CRT_MEGA_SMEM_STACK
#define CRT_MEGA_SMEM_STACK 0
#endif
template <bool COUNT>
__global__ void __launch_bounds__(256) k_render_mega(RenderArgs a) {
#if CRT_MEGA_SMEM_STACK > 0
    __shared__ uint2 sstack[CRT_MEGA_SMEM_STACK][256];
    TravStackShared<CRT_MEGA_SMEM_STACK, 256> st;
    st.slice = &sstack[0][threadIdx.x];
#else
    TravStack st;
#endif
    int x, y;
    const bool valid = tile_pixel(a.rc.width, a.rc.height, x, y);
    unsigned long long rays = 0, nodes = 0, prims = 0;
    if (valid) {
        const size_t n = (size_t)a.rc.width * a.rc.height;
        const size_t pix = (size_t)y * a.rc.width + x;
        Rng rng = load_rng(a.fb.rng, n, pix);
        vec3 sum = mk3(0.f, 0.f, 0.f);
        float nprev = 0.f;
        if (a.flags & CRT_RENDER_ACCUMULATE) {
            const float4 acc = a.fb.accum[pix];
            sum = mk3(acc.x, acc.y, acc.z);
            nprev = acc.w;
        }
        TraceCounters ctr;
        ctr.nodes = 0; ctr.prims = 0;
        vec3 o = mk3(0.f, 0.f, 0.f), d = mk3(0.f, 0.f, 1.f), thr = mk3(1.f, 1.f, 1.f);
        int sample = 0, bounce = 0;
        bool alive = false;
        uint32_t nrays = 0;
        while (true) {
            if (!alive) {
                if (sample == a.rc.spp) break;
                camera_ray(a.cam, x, y, a.rc.width, a.rc.height, rng, o, d);
                thr = mk3(1.f, 1.f, 1.f);
                bounce = 0;
                alive = true;
            }
            if (bounce >= a.rc.min_bounces) {             // Russian roulette, CUDAKernels.h:113-121
                const float p = fminf(fmaxf(thr.x, fmaxf(thr.y, thr.z)), a.rc.rr_max_prob);
                if (rng_uniform(rng) > p) { alive = false; sample++; continue; }
                thr = v_scale(thr, f_rcp(p));
            }
            HitRec h;
            nrays++;
            if (trace_closest_on<COUNT>(st, a.scene, o, d, 0.001f, INFINITY, h, &ctr)) {
                const Surface s = surface_at(a.scene, h, o, d);
                if (s.mat < a.scene.num_mats) {
                    const MatRec m = a.scene.mats[s.mat];
                    vec3 att, no, nd;
                    if (scatter(m, s, d, rng, att, no, nd)) {
                        thr = v_mul(thr, att);
                        o = no;
                        d = nd;
                    } else {                               // raw emission ends the path, :132-134
                        sum = v_add(sum, emitted(m));
                        alive = false;
                        sample++;
                        continue;
                    }
                }
            } else {                                       // sky, :137-142
                sum = v_add(sum, v_mul(thr, background(d)));
                alive = false;
                sample++;
                continue;
            }
            bounce++;
            if (bounce >= a.rc.max_bounces) { alive = false; sample++; }   // exhausted: black
        }
        const float ntot = nprev + (float)a.rc.spp;
        a.fb.accum[pix] = make_float4(sum.x, sum.y, sum.z, ntot);
        if (!(a.flags & CRT_RENDER_NO_TONEMAP)) {
            const float scale = f_div(1.f, ntot);          // m_PixelSampleScale, Camera.cuh:71
            a.fb.rgba8[pix] = pack_rgba8(v_scale(sum, scale));
        }
        store_rng(a.fb.rng, n, pix, rng);
        rays = nrays;
        nodes = ctr.nodes;
        prims = ctr.prims;
    }
    rays = warp_sum(rays);
    if (COUNT) { nodes = warp_sum(nodes); prims = warp_sum(prims); }
    if ((threadIdx.x & 31) == 0) {
        if (rays) atomicAdd(&a.fb.counters[0], rays);
        if (COUNT) { atomicAdd(&a.fb.counters[2], nodes); atomicAdd(&a.fb.counters[3], prims); }
    }
}

cudaError_t launch_render_mega(const RenderArgs& a, bool counted, int num_sms, cudaStream_t stream, uint32_t* launches) {
    (void)num_sms;
    const int tiles = ((a.rc.width + 15) / 16) * ((a.rc.height + 15) / 16);
    if (counted) k_render_mega<true><<<tiles, 256, 0, stream>>>(a);
    else k_render_mega<false><<<tiles, 256, 0, stream>>>(a);
    if (launches) *launches += 1;
    return cudaGetLastError();
}

// ---- primary-hit ids (parity buffer) ----------------------------------------------------------
__global__ void __launch_bounds__(256) k_primary(SceneView scene, Camera cam, int width, int height, unsigned long long seed,
                                                 unsigned long long base, int32_t* ids, float* tout) {
    int x, y;
    if (!tile_pixel(width, height, x, y)) return;
    const size_t pix = (size_t)y * width + x;
    curandStateXORWOW_t st;
    curand_init(seed, base + pix, 0, &st);
    Rng rng;
    rng.v0 = st.v[0]; rng.v1 = st.v[1]; rng.v2 = st.v[2]; rng.v3 = st.v[3]; rng.v4 = st.v[4]; rng.d = st.d;
    vec3 o, d;
    camera_ray(cam, x, y, width, height, rng, o, d);
    HitRec h;
    const bool hit = trace_closest<false>(scene, o, d, 0.001f, INFINITY, h, nullptr);
    ids[pix] = hit ? ((scene.key_to_id && (uint32_t)h.id < scene.num_tris) ? (int32_t)scene.key_to_id[h.id] : h.id) : -1;
    if (tout) tout[pix] = hit ? h.t : INFINITY;
}

cudaError_t launch_primary(const SceneView& scene, const Camera& cam, int width, int height, unsigned long long seed,
                           unsigned long long subseq_base, int32_t* ids, float* t, cudaStream_t stream) {
    const int tiles = ((width + 15) / 16) * ((height + 15) / 16);
    k_primary<<<tiles, 256, 0, stream>>>(scene, cam, width, height, seed, subseq_base, ids, t);
    return cudaGetLastError();
}

// ---- primary-hit ids without any allocation (crt_primary_ids_static) --------------------------
// The id / t staging buffers that crt_primary_ids takes from cudaMalloc live here in a file-scope device array
// instead: CRT_PRIMARY_STATIC_MAX_PIXELS * 8 bytes = 8 MB, held by every GPU that loads the library until the
// process exits. In host code `g_primary_scratch` is only a stand-in for the symbol, not a device address: the
// address each device maps it at is looked up once per device (that device current) with cudaGetSymbolAddress
// and handed to the kernel as an ordinary pointer. Every call on a device shares the array: k_primary writes
// each of the frame's pixels (nothing has to start at zero, stale contents past width*height are never read),
// and two calls that overlap on the same GPU race on it.
static __device__ __align__(256) unsigned char g_primary_scratch[(size_t)CRT_PRIMARY_STATIC_MAX_PIXELS * 8];

static constexpr int kMaxStaticDevices = 64;
static std::mutex g_primary_scratch_mutex;
static unsigned char* g_primary_scratch_ptr[kMaxStaticDevices] = {};

cudaError_t launch_primary_static(const SceneView& scene, const Camera& cam, int width, int height, unsigned long long seed,
                                  unsigned long long subseq_base, bool want_t, int device, cudaStream_t stream,
                                  int32_t** d_ids, float** d_t) {
    const size_t n = (size_t)width * height;
    if (device < 0 || device >= kMaxStaticDevices || n == 0 || n > (size_t)CRT_PRIMARY_STATIC_MAX_PIXELS) return cudaErrorInvalidValue;
    unsigned char* base = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_primary_scratch_mutex);
        if (!g_primary_scratch_ptr[device]) {
            int current = -1;
            cudaError_t e = cudaGetDevice(&current);
            if (e != cudaSuccess) return e;
            if (current != device) return cudaErrorInvalidDevice;   // the caller makes the context's device current first
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_primary_scratch);
            if (e != cudaSuccess) return e;
            if (!p) return cudaErrorInvalidSymbol;
            g_primary_scratch_ptr[device] = static_cast<unsigned char*>(p);
        }
        base = g_primary_scratch_ptr[device];
    }
    int32_t* ids = reinterpret_cast<int32_t*>(base);
    float* t = want_t ? reinterpret_cast<float*>(base + (size_t)CRT_PRIMARY_STATIC_MAX_PIXELS * 4) : nullptr;
    const int tiles = ((width + 15) / 16) * ((height + 15) / 16);
    k_primary<<<tiles, 256, 0, stream>>>(scene, cam, width, height, seed, subseq_base, ids, t);
    *d_ids = ids;
    *d_t = t;
    return cudaGetLastError();
}

// ---- tonemap: writeColor over the accumulation buffer (CRTUtility.cuh:21-32) -------------------
__global__ void __launch_bounds__(256) k_tonemap(const float4* accum, uint32_t* rgba8, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = accum[i];
        const float scale = f_div(1.f, a.w);
        rgba8[i] = a.w > 0.f ? pack_rgba8(v_scale(mk3(a.x, a.y, a.z), scale)) : 0xff000000u;
    }
}

cudaError_t launch_tonemap(const FrameBuffers& fb, int width, int height, cudaStream_t stream) {
    const size_t n = (size_t)width * height;
    int blocks = (int)((n + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    k_tonemap<<<blocks, 256, 0, stream>>>(fb.accum, fb.rgba8, n);
    return cudaGetLastError();
}

}  // namespace crt
