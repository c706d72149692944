// l2_bw.cu — L2-resident streaming-read bandwidth of the GPU (the bound of the cache-resident
// scenes is L2, not HBM). Every CTA of a 148 x 8 grid reads the whole working set `reps` times with 16-byte
// ld.global.cg loads (L1 bypassed), so the bytes counted all cross the L2 -> SM path.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o raytracer-cuda_b200/build/l2_bw scripts/l2_bw.cu
//   raytracer-cuda_b200/build/l2_bw > gpurun_out/l2_peak.json
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

__global__ void __launch_bounds__(256) k_read(const uint4* p, size_t n, int reps, unsigned* sink) {
    unsigned acc = 0;
    for (int r = 0; r < reps; r++) {
        // CTAs start at different offsets so that they do not all hit the same L2 slice at once
        const size_t start = ((size_t)blockIdx.x * 4099u * 256u) % n;
        for (size_t i = threadIdx.x; i < n; i += 256 * 4) {
            uint4 v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                size_t j = start + i + (size_t)k * 256;
                if (j >= n) j -= n;
                if (j >= n) j -= n;
                asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[k].x), "=r"(v[k].y), "=r"(v[k].z), "=r"(v[k].w) : "l"(p + j));
            }
#pragma unroll
            for (int k = 0; k < 4; k++) acc += v[k].x ^ v[k].y ^ v[k].z ^ v[k].w;
        }
    }
    if (acc == 0x12345678u) *sink = acc;
}

int main() {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int grid = prop.multiProcessorCount * 8;
    unsigned* sink;
    cudaMalloc(&sink, 4);
    const size_t sizes_mb[] = {4, 16, 64, 512};
    printf("{\"device\": \"%s\", \"sms\": %d, \"l2_bytes\": %d, \"how\": \"148x8 CTAs of 256, every CTA streams the whole working set with ld.global.cg.v4 (L1 bypassed), CUDA events, best of 5\", \"points\": [", prop.name, prop.multiProcessorCount, prop.l2CacheSize);
    for (int s = 0; s < 4; s++) {
        const size_t bytes = sizes_mb[s] << 20, n = bytes / 16;
        uint4* p;
        cudaMalloc(&p, bytes);
        cudaMemset(p, 1, bytes);
        const int reps = sizes_mb[s] <= 64 ? (int)(64 / sizes_mb[s]) : 1;   // 64 MB read per CTA and launch (512 MB: once)
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        float best = 1e30f;
        for (int it = 0; it < 6; it++) {
            cudaEventRecord(e0);
            k_read<<<grid, 256>>>(p, n, reps, sink);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (it > 0 && ms < best) best = ms;
        }
        const double gbs = (double)bytes * reps * grid / (best * 1e-3) / 1e9;
        printf("%s{\"working_set_mb\": %zu, \"ms\": %.3f, \"gbs\": %.1f}", s ? ", " : "", sizes_mb[s], best, gbs);
        cudaFree(p);
    }
    printf("], \"error\": \"%s\"}\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
