// crt_internal.h — declarations shared by the library's translation units (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include "crt_bvh_driver.h"

namespace crt {

// crt_bvh_build.cu
int build_bvh_device(const SceneInput& dev_in, uint32_t leaf_max, cudaStream_t stream, int num_sms, WideBvh& out,
                     float* build_ms, cudaError_t* cuda_err);

// Per-context device buffers handed to the kernels.
struct FrameBuffers {
    uint32_t* rng;        // 6 planes of width*height: v0..v4, d
    float4* accum;        // {sum_r, sum_g, sum_b, n_samples}
    uint32_t* rgba8;      // packed RGBA8, row 0 = bottom
    unsigned long long* counters;   // [0] rays [1] samples [2] nodes fetched [3] prims tested
};

struct RenderArgs {
    SceneView scene;
    Camera cam;
    RenderConst rc;
    FrameBuffers fb;
    uint32_t flags;       // CRT_RENDER_*
};

struct KernelTimes {
    float extend_ms = 0.f, shade_ms = 0.f, tail_ms = 0.f, begin_ms = 0.f, compact_ms = 0.f;
    uint32_t extend_launches = 0, shade_launches = 0;
    uint32_t host_syncs = 0, compactions = 0, iterations = 0, graph_launches = 0;   // blocking stream synchronisations / queue compactions / wavefront iterations of the frame
};

// crt_render.cu — every launcher enqueues on `stream` and returns the launch error (no sync).
cudaError_t launch_rng_init(uint32_t* rng_planes, int width, int height, unsigned long long seed,
                            unsigned long long subseq_base, int num_sms, cudaStream_t stream);
cudaError_t launch_render_mega(const RenderArgs& a, bool counted, int num_sms, cudaStream_t stream, uint32_t* launches);
cudaError_t launch_primary(const SceneView& scene, const Camera& cam, int width, int height, unsigned long long seed,
                           unsigned long long subseq_base, int32_t* ids, float* t, cudaStream_t stream);
// Same kernel, ids / t staged in the library's file-scope device array (no allocation); returns where they are.
cudaError_t launch_primary_static(const SceneView& scene, const Camera& cam, int width, int height, unsigned long long seed,
                                  unsigned long long subseq_base, bool want_t, int device, cudaStream_t stream,
                                  int32_t** d_ids, float** d_t);
cudaError_t launch_tonemap(const FrameBuffers& fb, int width, int height, cudaStream_t stream);

// crt_wavefront.cu
struct WavefrontState;   // persistent path state, owned by the context
cudaError_t wavefront_create(WavefrontState** out, int width, int height);
void wavefront_destroy(WavefrontState* s);
cudaError_t launch_render_wavefront(WavefrontState* ws, const RenderArgs& a, bool counted, int num_sms, cudaStream_t stream,
                                    uint32_t* launches, KernelTimes* times, bool* incomplete);

}  // namespace crt
