// crt_bvh_build.cu — CUDA execution policy for the BVH build (sm_100a).
//
// One kernel launch per build step (grid-stride over the items, grid sized in multiples of the SM
// count), cub::DeviceRadixSort for the Morton order, cub::DeviceScan for the PLOC compaction.
// Replaces the reference's <<<1,1>>> builders (reference: CudaRayTracer/src/CUDAKernels.h:86-100).
#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <chrono>
#include <cstdio>
#include <cstdlib>

#include "crt_bvh_driver.h"
#include "crt_internal.h"

namespace crt {

template <class F>
__global__ void __launch_bounds__(256) k_for_n(int n, F f) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) f(i);
}

// All remaining PLOC rounds of at most CRT_PLOC_SMALL clusters in one CTA (crt_bvh_driver.h, PlocSmallArgs): the same
// functors as the per-round launches, __syncthreads() between the steps (it orders the block's global-memory accesses),
// and a block-wide exclusive scan (two items per thread) in place of cub::DeviceScan.
static_assert(CRT_PLOC_SMALL <= 2048, "k_ploc_small scans two items per thread of a 1024-thread CTA");
__global__ void __launch_bounds__(1024) k_ploc_small(PlocSmallArgs a) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_total;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t* cl_in = a.cl_in;
    uint32_t* cl_out = a.cl_out;
    int n = a.n;
    uint32_t rounds = 0;
    while (n > 1) {
        PlocNearestFn fn;
        fn.blo = a.blo; fn.bhi = a.bhi; fn.cluster = cl_in; fn.nearest = a.nearest; fn.n = n;
        for (int i = tid; i < n; i += 1024) fn(i);
        __syncthreads();
        PlocMergeFn fm;
        fm.blo = a.blo; fm.bhi = a.bhi; fm.bcount = a.bcount; fm.cluster = cl_in; fm.nearest = a.nearest; fm.merged = a.merged;
        fm.flags = a.flags; fm.node_counter = a.node_counter;
        for (int i = tid; i < n; i += 1024) fm(i);
        __syncthreads();
        const uint32_t f0 = 2 * tid < n ? a.flags[2 * tid] : 0u, f1 = 2 * tid + 1 < n ? a.flags[2 * tid + 1] : 0u;
        const uint32_t mine = f0 + f1;
        uint32_t incl = mine;
        __syncwarp();   // the loops above leave the lanes of a warp with different trip counts: explicit rendezvous
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const uint32_t w = s_warp[lane];
            uint32_t wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, wi, d);
                if (lane >= d) wi += v;
            }
            s_warp[lane] = wi - w;
            if (lane == 31) s_total = wi;
        }
        __syncthreads();
        const uint32_t excl = s_warp[warp] + incl - mine;
        if (2 * tid < n) a.pos[2 * tid] = excl;
        if (2 * tid + 1 < n) a.pos[2 * tid + 1] = excl + f0;
        const uint32_t total = s_total;
        __syncthreads();
        PlocCompactFn fc;
        fc.merged = a.merged; fc.flags = a.flags; fc.pos = a.pos; fc.cluster_out = cl_out; fc.total = a.total; fc.n = n;
        for (int i = tid; i < n; i += 1024) fc(i);
        __syncthreads();
        if ((int)total >= n || total == 0u) {   // no progress: cannot happen (see PlocNearestFn)
            if (tid == 0) { a.result[0] = 0xffffffffu; a.result[1] = 0u; }
            return;
        }
        n = (int)total;
        uint32_t* t = cl_in; cl_in = cl_out; cl_out = t;
        rounds++;
    }
    if (tid == 0) { a.result[0] = rounds; a.result[1] = cl_in[0]; }
}

struct CudaExec {
    cudaStream_t stream;
    int num_sms;
    cudaError_t err = cudaSuccess;
    void* cub_tmp = nullptr;
    size_t cub_tmp_bytes = 0;

    void note(cudaError_t e) { if (err == cudaSuccess && e != cudaSuccess) err = e; }

    double alloc_host_ms = 0.0, free_host_ms = 0.0;   // host time spent inside cudaMalloc / cudaFree (CRT_BUILD_TRACE)
    size_t alloc_bytes = 0;
    template <class T> T* alloc(size_t n) {
        void* p = nullptr;
        const auto t0 = std::chrono::steady_clock::now();
        cudaError_t e = cudaMalloc(&p, (n ? n : 1) * sizeof(T));
        alloc_host_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        alloc_bytes += (n ? n : 1) * sizeof(T);
        if (e != cudaSuccess) { note(e); return nullptr; }
        return (T*)p;
    }
    void release(void* p) {
        if (!p) return;
        const auto t0 = std::chrono::steady_clock::now();
        note(cudaFree(p));
        free_host_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }

    // Temporaries of one build come out of ONE arena (bump allocation, 256-byte aligned, released as a whole by
    // free_arena): a build used to spend more host time inside ~30 cudaMalloc / cudaFree calls (each cudaFree also
    // drains the device) than the GPU spent on its kernels. A request the arena cannot hold falls back to cudaMalloc.
    char* arena = nullptr;
    size_t arena_bytes = 0, arena_used = 0;
    void reserve_tmp(size_t bytes) {
        if (arena || getenv("CRT_BUILD_NO_ARENA")) return;
        arena = alloc<char>(bytes);
        arena_bytes = arena ? bytes : 0;
        arena_used = 0;
    }
    template <class T> T* alloc_tmp(size_t n) {
        const size_t bytes = ((n ? n : 1) * sizeof(T) + 255) & ~(size_t)255;
        if (arena && arena_used + bytes <= arena_bytes) {
            T* p = reinterpret_cast<T*>(arena + arena_used);
            arena_used += bytes;
            return p;
        }
        return alloc<T>(n);
    }
    void release_tmp(void* p) {
        if (arena && (char*)p >= arena && (char*)p < arena + arena_bytes) return;
        release(p);
    }
    void free_arena() {
        if (arena) release(arena);
        arena = nullptr;
        arena_bytes = arena_used = 0;
    }

    template <class F> void for_n(int n, const F& f) {
        if (n <= 0) return;
        int blocks = (n + 255) / 256;
        const int cap = num_sms * 8;             // 8 resident 256-thread CTAs per SM
        if (blocks > cap) blocks = cap;
        k_for_n<F><<<blocks, 256, 0, stream>>>(n, f);
        note(cudaGetLastError());
    }

    void ensure_tmp(size_t bytes) {
        if (bytes <= cub_tmp_bytes) return;
        if (cub_tmp) cudaFree(cub_tmp);
        cub_tmp = nullptr;
        cub_tmp_bytes = 0;
        if (cudaMalloc(&cub_tmp, bytes) == cudaSuccess) cub_tmp_bytes = bytes;
        else note(cudaErrorMemoryAllocation);
    }

    void sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t* keys_tmp, uint32_t* vals_tmp, int n, uint64_t** ks, uint32_t** vs) {
        size_t bytes = 0;
        note(cub::DeviceRadixSort::SortPairs(nullptr, bytes, keys, keys_tmp, vals, vals_tmp, n, 0, 63, stream));
        ensure_tmp(bytes);
        if (cub_tmp) note(cub::DeviceRadixSort::SortPairs(cub_tmp, bytes, keys, keys_tmp, vals, vals_tmp, n, 0, 63, stream));
        *ks = keys_tmp;
        *vs = vals_tmp;
    }

    void exclusive_scan(const uint32_t* in, uint32_t* out, int n) {
        size_t bytes = 0;
        note(cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, n, stream));
        ensure_tmp(bytes);
        if (cub_tmp) note(cub::DeviceScan::ExclusiveSum(cub_tmp, bytes, in, out, n, stream));
    }

    void ploc_small(const PlocSmallArgs& a) {
        k_ploc_small<<<1, 1024, 0, stream>>>(a);
        note(cudaGetLastError());
    }

    void fill_u32(uint32_t* p, uint32_t v, size_t n) {
        if (v == 0u) note(cudaMemsetAsync(p, 0, n * sizeof(uint32_t), stream));
        else note(cudaErrorInvalidValue);   // only zero fills are needed
    }

    void upload(void* d, const void* s, size_t b) {
        note(cudaMemcpyAsync(d, s, b, cudaMemcpyHostToDevice, stream));
        note(cudaStreamSynchronize(stream));   // the source is usually a stack temporary
    }
    void download(void* d, const void* s, size_t b) {
        note(cudaMemcpyAsync(d, s, b, cudaMemcpyDeviceToHost, stream));
        note(cudaStreamSynchronize(stream));
    }
};

int build_bvh_device(const SceneInput& dev_in, uint32_t leaf_max, cudaStream_t stream, int num_sms, WideBvh& out,
                     float* build_ms, cudaError_t* cuda_err) {
    CudaExec ex;
    ex.stream = stream;
    ex.num_sms = num_sms > 0 ? num_sms : 148;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, stream);
    int rc = build_wide_bvh(ex, dev_in, leaf_max, out);
    cudaEventRecord(e1, stream);
    ex.note(cudaStreamSynchronize(stream));
    ex.free_arena();   // after the stream has drained: nothing of the build still reads it
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, e0, e1) != cudaSuccess) ms = 0.f;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (ex.cub_tmp) cudaFree(ex.cub_tmp);
    if (getenv("CRT_BUILD_TRACE"))
        fprintf(stderr, "bvh_build_trace: %.2f ms between the stream events; host time inside cudaMalloc %.2f ms (%zu MB in temporaries + outputs), inside cudaFree %.2f ms\n",
                ms, ex.alloc_host_ms, ex.alloc_bytes >> 20, ex.free_host_ms);
    if (build_ms) *build_ms = ms;
    if (cuda_err) *cuda_err = ex.err;
    if (ex.err != cudaSuccess) return CRT_E_CUDA;
    return rc;
}

}  // namespace crt
