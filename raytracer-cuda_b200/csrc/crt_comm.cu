// crt_comm.cu — multi-GPU behind the C-ABI: the spp budget of a frame sharded over GPUs, ONE sum-reduce of the
// float4 accumulation buffer onto a root, tonemap there.
//
// The reference is single-GPU (one device, default stream: CudaRayTracer/src/CUDARenderer.cuh:55-60; no
// cudaSetDevice / NCCL anywhere under CudaRayTracer/src). What shards is its sample loop
// (src/CUDAKernels.h:158-162): every (pixel, sample) is independent, so GPU g renders its share of the samples
// of EVERY pixel from the cuRAND subsequences g*W*H + pixel and the per-pixel sums add up.
//
// Two ways in:
//  * one process per GPU (torchrun / mpirun): crt_comm_unique_id on rank 0, the caller hands the 128 bytes to the
//    other ranks by whatever means it has, crt_comm_init on every rank, crt_reduce_accum after crt_render.
//  * one process, several GPUs: crt_group_* keeps one context per device, renders the shards on one host thread
//    per device and reduces either with ONE kernel on the root GPU that reads the peers' buffers over NVLink
//    (peer access), adds them and tonemaps in the same pass, or, without peer access, with ncclReduce.
// NCCL is bound at run time (dlopen of libnccl.so.2: the copy the process already holds, e.g. PyTorch's, or the
// system's), so the library has no link-time dependency on it and single-GPU use never touches it.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "crt_ctx.h"
#include "crt_shade.h"

using namespace crt;

namespace {

// the handful of NCCL entry points used, with the types of nccl.h (2.x ABI)
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId_ { char internal[128]; };
enum { ncclSuccess_ = 0, ncclFloat_ = 7, ncclSum_ = 0 };
struct NcclApi {
    void* so = nullptr;
    int (*GetUniqueId)(ncclUniqueId_*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId_, int) = nullptr;
    int (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    std::string why;
};

NcclApi* nccl() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return &api;
    tried = true;
    // 1. a copy of NCCL the process already holds (PyTorch brings its own libnccl.so.2): two NCCL versions cannot share
    //    the soname, so that one is used. 2. CRT_NCCL_LIB (a path). 3. the system's libnccl.so.2.
    // A process that will load PyTorch LATER should import it before the first crt_comm_* / crt_group_* call (or set
    // CRT_NCCL_LIB to PyTorch's copy): otherwise PyTorch would find the system's older NCCL under its soname.
    api.so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!api.so) if (const char* path = getenv("CRT_NCCL_LIB")) api.so = dlopen(path, RTLD_NOW | RTLD_LOCAL);
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        if (api.so) break;
        api.so = dlopen(n, RTLD_NOW | RTLD_LOCAL);
    }
    if (!api.so) { api.why = std::string("NCCL not found (dlopen libnccl.so.2): ") + (dlerror() ? dlerror() : ""); return &api; }
    bool ok = true;
    auto sym = [&](const char* name) { void* p = dlsym(api.so, name); if (!p) { ok = false; api.why = std::string("NCCL symbol missing: ") + name; } return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.Reduce = (decltype(api.Reduce))sym("ncclReduce");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
    if (!ok) { dlclose(api.so); api.so = nullptr; }
    return &api;
}

int fail(crt_ctx* ctx, int code, const std::string& msg) { return crt_ctx_fail_(ctx, code, msg.c_str()); }
int fail_nccl(crt_ctx* ctx, int rc, const char* what) {
    NcclApi* a = nccl();
    return fail(ctx, CRT_E_CUDA, std::string(what) + ": NCCL error " + std::to_string(rc) + " (" + (a->GetErrorString ? a->GetErrorString(rc) : "?") + ")");
}

// ---- reduce + tonemap in one pass over peer memory ------------------------------------------------------
// Root GPU: accum[i] = sum over the G buffers (its own and the peers', read through NVLink), rgba8[i] = writeColor
// of the mean (reference: CRTUtility.cuh:21-32). Fixed summation order 0..G-1, so the result does not depend on
// the collective's algorithm. HBM/NVLink bound: 16 B x G read + 20 B written per pixel.
struct PeerBuffers {
    const float4* buf[CRT_GROUP_MAX_DEVICES];
    int count;
};
__global__ void __launch_bounds__(256) k_reduce_tonemap(PeerBuffers peers, float4* accum, uint32_t* rgba8, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float4 s = __ldcs(peers.buf[0] + i);
        for (int g = 1; g < peers.count; g++) {
            const float4 v = __ldcs(peers.buf[g] + i);
            s.x = f_add(s.x, v.x); s.y = f_add(s.y, v.y); s.z = f_add(s.z, v.z); s.w = f_add(s.w, v.w);
        }
        accum[i] = s;
        rgba8[i] = s.w > 0.f ? pack_rgba8(v_scale(mk3(s.x, s.y, s.z), f_div(1.f, s.w))) : 0xff000000u;
    }
}

}  // namespace

void crt_comm_release_(crt_ctx* ctx) {
    if (ctx && ctx->comm) {
        NcclApi* a = nccl();
        if (a->so) a->CommDestroy((ncclComm_t)ctx->comm);
        ctx->comm = nullptr;
        ctx->comm_nranks = 1;
        ctx->comm_rank = 0;
    }
}

struct crt_group {
    std::vector<crt_ctx*> ctx;
    std::vector<int> device;
    bool peer = false;            // device[0] can read every other device's memory
    bool nccl_ready = false;      // communicators created (lazily, only without peer access or when forced)
    int width = 0, height = 0;
    float reduce_ms = 0.f;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string err;
};

// run f(i) for every device on its own host thread (crt_* calls block); first failure wins
template <class F>
static int group_each(crt_group* g, const char* what, F f) {
    const int n = (int)g->ctx.size();
    std::vector<int> rc(n, CRT_OK);
    try {
        std::vector<std::thread> th;
        for (int i = 1; i < n; i++) th.emplace_back([&, i] { rc[i] = f(i); });
        rc[0] = f(0);
        for (std::thread& t : th) t.join();
    } catch (...) {
        g->err = std::string(what) + ": could not start the device threads";
        return CRT_E_NOMEM;
    }
    for (int i = 0; i < n; i++)
        if (rc[i] != CRT_OK) { g->err = std::string(what) + " (device " + std::to_string(g->device[i]) + "): " + crt_last_error(g->ctx[i]); return rc[i]; }
    return CRT_OK;
}

static int group_nccl_init(crt_group* g) {
    if (g->nccl_ready) return CRT_OK;
    NcclApi* a = nccl();
    if (!a->so) { g->err = "crt_group_render: no peer access between the devices and " + a->why; return CRT_E_INVALID; }
    std::vector<ncclComm_t> comms(g->ctx.size());
    const int rc = a->CommInitAll(comms.data(), (int)g->ctx.size(), g->device.data());
    if (rc != ncclSuccess_) { g->err = std::string("ncclCommInitAll: ") + a->GetErrorString(rc); return CRT_E_CUDA; }
    for (size_t i = 0; i < g->ctx.size(); i++) { g->ctx[i]->comm = comms[i]; g->ctx[i]->comm_nranks = (int)g->ctx.size(); g->ctx[i]->comm_rank = (int)i; }
    g->nccl_ready = true;
    return CRT_OK;
}

extern "C" {

int crt_comm_unique_id(void* out_id, size_t bytes) {
    if (!out_id || bytes < CRT_COMM_ID_BYTES) return fail(nullptr, CRT_E_INVALID, "crt_comm_unique_id: need a 128-byte buffer");
    NcclApi* a = nccl();
    if (!a->so) return fail(nullptr, CRT_E_INVALID, "crt_comm_unique_id: " + a->why);
    ncclUniqueId_ id;
    const int rc = a->GetUniqueId(&id);
    if (rc != ncclSuccess_) return fail_nccl(nullptr, rc, "ncclGetUniqueId");
    memcpy(out_id, &id, CRT_COMM_ID_BYTES);
    return CRT_OK;
}

int crt_comm_init(crt_ctx* ctx, const void* id, int32_t nranks, int32_t rank) {
    if (!ctx) return CRT_E_INVALID;
    if (!id || nranks < 1 || rank < 0 || rank >= nranks) return fail(ctx, CRT_E_INVALID, "crt_comm_init: bad arguments");
    NcclApi* a = nccl();
    if (!a->so) return fail(ctx, CRT_E_INVALID, "crt_comm_init: " + a->why);
    crt_comm_release_(ctx);
    cudaError_t e = cudaSetDevice(ctx->cfg.device);
    if (e != cudaSuccess) return fail(ctx, CRT_E_CUDA, std::string("crt_comm_init: cudaSetDevice: ") + cudaGetErrorString(e));
    ncclUniqueId_ uid;
    memcpy(&uid, id, CRT_COMM_ID_BYTES);
    ncclComm_t comm = nullptr;
    const int rc = a->CommInitRank(&comm, nranks, uid, rank);
    if (rc != ncclSuccess_) return fail_nccl(ctx, rc, "ncclCommInitRank");
    ctx->comm = comm;
    ctx->comm_nranks = nranks;
    ctx->comm_rank = rank;
    return CRT_OK;
}

int crt_comm_destroy(crt_ctx* ctx) {
    if (!ctx) return CRT_E_INVALID;
    cudaSetDevice(ctx->cfg.device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    crt_comm_release_(ctx);
    return CRT_OK;
}

int crt_comm_info(crt_ctx* ctx, int32_t* nranks, int32_t* rank, int32_t* nccl_version) {
    if (!ctx) return CRT_E_INVALID;
    if (nranks) *nranks = ctx->comm_nranks;
    if (rank) *rank = ctx->comm_rank;
    if (nccl_version) { *nccl_version = 0; NcclApi* a = nccl(); if (a->so) { int v = 0; if (a->GetVersion(&v) == ncclSuccess_) *nccl_version = v; } }
    return CRT_OK;
}

int crt_reduce_accum(crt_ctx* ctx, int32_t root) {
    if (!ctx) return CRT_E_INVALID;
    if (!ctx->comm || ctx->comm_nranks == 1) {                 // single GPU: the "reduce" is the identity, root tonemaps
        if (root != 0) return fail(ctx, CRT_E_INVALID, "crt_reduce_accum: bad root");
        return crt_tonemap(ctx);
    }
    if (root < 0 || root >= ctx->comm_nranks) return fail(ctx, CRT_E_INVALID, "crt_reduce_accum: bad root");
    NcclApi* a = nccl();
    cudaError_t e = cudaSetDevice(ctx->cfg.device);
    if (e != cudaSuccess) return fail(ctx, CRT_E_CUDA, std::string("crt_reduce_accum: cudaSetDevice: ") + cudaGetErrorString(e));
    const size_t count = (size_t)ctx->cfg.width * ctx->cfg.height * 4;
    // in place, on the stream the render ran on: ordered behind this rank's last kernel and before the root's tonemap
    const int rc = a->Reduce(ctx->fb.accum, ctx->fb.accum, count, ncclFloat_, ncclSum_, root, (ncclComm_t)ctx->comm, ctx->stream);
    if (rc != ncclSuccess_) return fail_nccl(ctx, rc, "ncclReduce");
    if (ctx->comm_rank == root) return crt_tonemap(ctx);       // launches on ctx->stream and synchronises it
    e = cudaStreamSynchronize(ctx->stream);                    // the buffer is free for the next frame on return
    if (e != cudaSuccess) return fail(ctx, CRT_E_CUDA, std::string("crt_reduce_accum: ") + cudaGetErrorString(e));
    return CRT_OK;
}

// ---- one process, several GPUs ------------------------------------------------------------------------------
const char* crt_group_last_error(const crt_group* g) { return g ? g->err.c_str() : crt_last_error(nullptr); }

void crt_group_destroy(crt_group* g) {
    if (!g) return;
    for (crt_ctx* c : g->ctx) crt_destroy(c);
    if (!g->device.empty()) cudaSetDevice(g->device[0]);
    if (g->ev0) cudaEventDestroy(g->ev0);
    if (g->ev1) cudaEventDestroy(g->ev1);
    delete g;
}

int crt_group_create(crt_group** out, const crt_config* cfg, const int32_t* devices, int32_t num_devices) {
    if (!out || !cfg || !devices || num_devices < 1 || num_devices > CRT_GROUP_MAX_DEVICES)
        return fail(nullptr, CRT_E_INVALID, "crt_group_create: need 1..CRT_GROUP_MAX_DEVICES devices");
    *out = nullptr;
    for (int i = 0; i < num_devices; i++)
        for (int j = 0; j < i; j++)
            if (devices[i] == devices[j]) return fail(nullptr, CRT_E_INVALID, "crt_group_create: a device is listed twice");
    crt_group* g = new (std::nothrow) crt_group();
    if (!g) return fail(nullptr, CRT_E_NOMEM, "crt_group_create: out of host memory");
    try {
        g->width = cfg->width; g->height = cfg->height;
        for (int i = 0; i < num_devices; i++) {
            crt_config c = *cfg;
            c.device = devices[i];
            crt_ctx* ctx = nullptr;
            const int rc = crt_create(&ctx, &c);
            if (rc != CRT_OK) { crt_group_destroy(g); return rc; }   // message already under the null context
            g->ctx.push_back(ctx);
            g->device.push_back(devices[i]);
        }
        // peer access root <- others: lets ONE kernel on the root read every shard (k_reduce_tonemap)
        g->peer = true;
        cudaSetDevice(devices[0]);
        for (int i = 1; i < num_devices && g->peer; i++) {
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, devices[0], devices[i]) != cudaSuccess || !can) { g->peer = false; break; }
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[i], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) g->peer = false;
            (void)cudaGetLastError();
        }
        cudaEventCreate(&g->ev0);
        cudaEventCreate(&g->ev1);
    } catch (...) {
        crt_group_destroy(g);
        return fail(nullptr, CRT_E_NOMEM, "crt_group_create: out of host memory");
    }
    *out = g;
    return CRT_OK;
}

int32_t crt_group_size(const crt_group* g) { return g ? (int32_t)g->ctx.size() : 0; }
crt_ctx* crt_group_ctx(crt_group* g, int32_t index) { return (g && index >= 0 && index < (int32_t)g->ctx.size()) ? g->ctx[index] : nullptr; }

int crt_group_set_scene(crt_group* g, const crt_scene* scene) {
    if (!g || !scene) return CRT_E_INVALID;
    return group_each(g, "crt_group_set_scene", [&](int i) { return crt_set_scene(g->ctx[i], scene); });   // BVH rebuilt per GPU: milliseconds
}
int crt_group_set_camera(crt_group* g, const crt_camera* cam) {
    if (!g || !cam) return CRT_E_INVALID;
    for (crt_ctx* c : g->ctx) { const int rc = crt_set_camera(c, cam); if (rc != CRT_OK) return rc; }
    return CRT_OK;
}
int crt_group_init_rng(crt_group* g, uint64_t seed) {
    if (!g) return CRT_E_INVALID;
    const uint64_t n = (uint64_t)g->width * (uint64_t)g->height;
    return group_each(g, "crt_group_init_rng", [&](int i) { return crt_init_rng(g->ctx[i], seed, (uint64_t)i * n); });
}

int crt_group_render(crt_group* g, const crt_render_params* params, uint32_t group_flags) {
    if (!g || !params) return CRT_E_INVALID;
    const int n = (int)g->ctx.size();
    if (params->spp < n) { g->err = "crt_group_render: fewer samples per pixel than devices"; return CRT_E_INVALID; }
    // shard g renders spp/G samples (+1 for the first spp % G shards) of every pixel, without tonemapping
    int rc = group_each(g, "crt_group_render", [&](int i) {
        crt_render_params p = *params;
        p.spp = params->spp / n + (i < params->spp % n ? 1 : 0);
        if (n > 1) p.flags |= CRT_RENDER_NO_TONEMAP;
        return crt_render(g->ctx[i], &p);
    });
    if (rc != CRT_OK || n == 1) return rc;
    crt_ctx* root = g->ctx[0];
    const size_t npix = (size_t)g->width * g->height;
    cudaSetDevice(g->device[0]);
    cudaEventRecord(g->ev0, root->stream);
    if (g->peer && !(group_flags & CRT_GROUP_FORCE_NCCL)) {
        // every shard is complete (crt_render synchronises): one pass on the root over its own and the peers' buffers
        PeerBuffers pb;
        pb.count = n;
        for (int i = 0; i < n; i++) pb.buf[i] = g->ctx[i]->fb.accum;
        k_reduce_tonemap<<<root->num_sms * 8, 256, 0, root->stream>>>(pb, root->fb.accum, root->fb.rgba8, npix);
        cudaEventRecord(g->ev1, root->stream);
        const cudaError_t e = cudaStreamSynchronize(root->stream);
        if (e != cudaSuccess) { g->err = std::string("k_reduce_tonemap: ") + cudaGetErrorString(e); (void)cudaGetLastError(); return CRT_E_CUDA; }
    } else {
        rc = group_nccl_init(g);
        if (rc != CRT_OK) return rc;
        NcclApi* a = nccl();
        a->GroupStart();
        int nrc = ncclSuccess_;
        for (int i = 0; i < n && nrc == ncclSuccess_; i++) {
            cudaSetDevice(g->device[i]);
            nrc = a->Reduce(g->ctx[i]->fb.accum, g->ctx[i]->fb.accum, npix * 4, ncclFloat_, ncclSum_, 0, (ncclComm_t)g->ctx[i]->comm, g->ctx[i]->stream);
        }
        const int erc = a->GroupEnd();
        if (nrc != ncclSuccess_ || erc != ncclSuccess_) { g->err = std::string("ncclReduce: ") + a->GetErrorString(nrc != ncclSuccess_ ? nrc : erc); return CRT_E_CUDA; }
        for (int i = 1; i < n; i++) { cudaSetDevice(g->device[i]); cudaStreamSynchronize(g->ctx[i]->stream); }
        cudaSetDevice(g->device[0]);
        rc = crt_tonemap(root);
        cudaEventRecord(g->ev1, root->stream);
        cudaStreamSynchronize(root->stream);
        if (rc != CRT_OK) { g->err = std::string("crt_tonemap: ") + crt_last_error(root); return rc; }
    }
    cudaEventElapsedTime(&g->reduce_ms, g->ev0, g->ev1);
    return CRT_OK;
}

int crt_group_get_stats(crt_group* g, crt_stats* out_sum, float* out_reduce_ms, int32_t* out_peer) {
    if (!g) return CRT_E_INVALID;
    if (out_sum) {
        crt_stats s = g->ctx[0]->stats;
        for (size_t i = 1; i < g->ctx.size(); i++) {
            const crt_stats& t = g->ctx[i]->stats;
            s.rays += t.rays; s.rays_total += t.rays_total; s.samples += t.samples; s.kernel_launches += t.kernel_launches;
            if (t.render_ms > s.render_ms) s.render_ms = t.render_ms;
        }
        *out_sum = s;
    }
    if (out_reduce_ms) *out_reduce_ms = g->reduce_ms;
    if (out_peer) *out_peer = g->peer ? 1 : 0;
    return CRT_OK;
}

}  // extern "C"
