// Replacement for CudaRayTracer/src/Core/Core.cuh used ONLY by the reference shim build.
// The upstream file does not compile with nvcc/g++ on Linux: its `swap` template names an
// undeclared `val2` (Core.cuh:17-23) and CudaAlloc::Memory calls checkCuda before it is
// declared (:61 vs :167); neither is used by the render path. What the render path does use
// is kept with identical meaning: the CUDA_KERNEL launch macro (:10) and swap_triplet (:25-39,
// used by Core/Mesh.cuh:187). <cstdint> is added for uint32_t (Core/HitInfo.cuh:10).
#pragma once
#include <cstdint>
#include <cmath>
#include <limits>

#define CUDA_KERNEL(...) <<< __VA_ARGS__ >>>

template <class T>
__host__ __device__ void swap_triplet(T& a1, T& a2, T& a3, T& b1, T& b2, T& b3) {
    T s1 = a1, s2 = a2, s3 = a3;
    a1 = b1; a2 = b2; a3 = b3;
    b1 = s1; b2 = s2; b3 = s3;
}
