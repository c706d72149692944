// crt_ctx.h — the context behind the opaque crt_ctx* of include/crt_api.h (shared by crt_api.cu and crt_comm.cu).
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <thread>

#include "../../include/crt_api.h"
#include "crt_internal.h"

struct crt_ctx {
    crt_config cfg{};
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t own_stream = nullptr;   // created by crt_create; `stream` may be replaced by crt_set_stream
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // frame buffers
    crt::FrameBuffers fb{};
    bool rng_ready = false;
    // scene
    float* d_pos = nullptr;
    uint32_t* d_idx = nullptr;
    int32_t* d_trimat = nullptr;
    float* d_sph = nullptr;
    int32_t* d_sphmat = nullptr;
    crt::MatRec* d_mats = nullptr;
    uint32_t* d_trikey = nullptr;    // reference-exact mode: tie key per triangle (its rank in the reference's order)
    uint32_t* d_key2id = nullptr;    // and the inverse, for the primary-id buffer
    uint32_t num_tris = 0, unseen_tris = 0;
    uint32_t num_mats = 0;
    crt::WideBvh bvh{};
    bool scene_ready = false;
    crt::Camera cam{};
    bool cam_ready = false;
    crt::WavefrontState* wf = nullptr;
    crt_stats stats{};
    // CRT_RENDER_ASYNC: the frame loop (which reads queue lengths back) runs on this host thread until crt_wait
    std::thread worker;
    bool async_pending = false;
    int async_rc = 0;
    crt_render_params async_params{};
    void* comm = nullptr;            // ncclComm_t of crt_comm_init (crt_comm.cu); null = single GPU
    int comm_rank = 0, comm_nranks = 1;
    std::string err;
};


// crt_api.cu
extern "C" int crt_ctx_fail_(crt_ctx* ctx, int code, const char* msg);
// crt_comm.cu
void crt_comm_release_(crt_ctx* ctx);   // called by crt_destroy
