// crt_api.cu — the C-ABI of include/crt_api.h: context, scene upload, camera, render, read-back.
//
// Host-side replacement for CUDARenderer + the device half of SceneManager
// (reference: CudaRayTracer/src/CUDARenderer.cuh:9-60, src/SceneManager.h:61-98).
// No CPU fallback: every entry point that needs the GPU fails with CRT_E_NO_DEVICE / CRT_E_CUDA.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/crt_api.h"
#include "crt_camera_host.h"
#include "crt_ctx.h"
#include "crt_internal.h"
#include "crt_trace.h"
#include "scene_gen.h"

using namespace crt;

static std::string g_create_error = "";

static int fail(crt_ctx* ctx, int code, const std::string& msg) {
    if (ctx) ctx->err = msg; else g_create_error = msg;
    return code;
}
static int fail_cuda(crt_ctx* ctx, cudaError_t e, const char* what) {
    (void)cudaGetLastError();
    return fail(ctx, CRT_E_CUDA, std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
}
#define CU(call)                                                        \
    do {                                                                \
        cudaError_t e__ = (call);                                       \
        if (e__ != cudaSuccess) return fail_cuda(ctx, e__, #call);      \
    } while (0)

// An asynchronous render (CRT_RENDER_ASYNC) owns the context until crt_wait: every other call is refused meanwhile.
#define NOT_BUSY(ctx)                                                                                              \
    do {                                                                                                           \
        if ((ctx) && (ctx)->async_pending) return CRT_E_BUSY;                                                     \
    } while (0)

static void free_scene(crt_ctx* ctx) {
    cudaFree(ctx->d_pos); cudaFree(ctx->d_idx); cudaFree(ctx->d_trimat); cudaFree(ctx->d_sph); cudaFree(ctx->d_sphmat);
    cudaFree(ctx->d_mats); cudaFree(ctx->bvh.nodes); cudaFree(ctx->bvh.prims); cudaFree(ctx->d_trikey); cudaFree(ctx->d_key2id);
    ctx->d_trikey = nullptr; ctx->d_key2id = nullptr;
    ctx->d_pos = nullptr; ctx->d_idx = nullptr; ctx->d_trimat = nullptr; ctx->d_sph = nullptr; ctx->d_sphmat = nullptr;
    ctx->d_mats = nullptr; ctx->bvh = WideBvh(); ctx->scene_ready = false;
}

static SceneView scene_view(const crt_ctx* ctx) {
    SceneView v;
    v.nodes = ctx->bvh.nodes; v.prims = ctx->bvh.prims; v.mats = ctx->d_mats;
    v.num_mats = ctx->num_mats; v.num_nodes = ctx->bvh.num_nodes; v.num_prims = ctx->bvh.num_prims;
    v.giant_base = ctx->bvh.giant_base; v.giant_count = ctx->bvh.giant_count;
    v.key_to_id = ctx->d_key2id; v.num_tris = ctx->num_tris;
    return v;
}

extern "C" {

int crt_api_version(void) { return CRT_API_VERSION; }

// internal helpers for crt_host.cpp / crt_obj.cpp (not declared in the public header)
int crt_ctx_size_(const crt_ctx* ctx, int* w, int* h) {
    if (!ctx) return CRT_E_INVALID;
    *w = ctx->cfg.width;
    *h = ctx->cfg.height;
    return CRT_OK;
}
int crt_ctx_fail_(crt_ctx* ctx, int code, const char* msg) { return fail(ctx, code, msg ? msg : ""); }

const char* crt_last_error(const crt_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int crt_create(crt_ctx** out_ctx, const crt_config* cfg) {
    crt_ctx* ctx = nullptr;   // for the CU macro: errors before allocation go to the global message
    if (!out_ctx || !cfg) return fail(nullptr, CRT_E_INVALID, "crt_create: null argument");
    *out_ctx = nullptr;
    if (cfg->width <= 0 || cfg->height <= 0 || (long long)cfg->width * cfg->height > (1ll << 30))
        return fail(nullptr, CRT_E_INVALID, "crt_create: bad image size");
    if (cfg->mode != CRT_MODE_WAVEFRONT && cfg->mode != CRT_MODE_MEGAKERNEL && cfg->mode != CRT_MODE_AUTO) return fail(nullptr, CRT_E_INVALID, "crt_create: bad mode");
    if (cfg->reference_exact != 0 && cfg->reference_exact != 1) return fail(nullptr, CRT_E_INVALID, "crt_create: reference_exact must be 0 or 1");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev <= 0) { (void)cudaGetLastError(); return fail(nullptr, CRT_E_NO_DEVICE, "crt_create: no CUDA device (this library has no CPU fallback)"); }
    if (cfg->device < 0 || cfg->device >= ndev) return fail(nullptr, CRT_E_INVALID, "crt_create: bad device ordinal");
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10) return fail(nullptr, CRT_E_NO_DEVICE, std::string("crt_create: kernels are built for sm_100a only, device is ") + prop.name);
    CU(cudaSetDevice(cfg->device));
    ctx = new (std::nothrow) crt_ctx();
    if (!ctx) return fail(nullptr, CRT_E_NOMEM, "crt_create: out of host memory");
    ctx->cfg = *cfg;
    if (ctx->cfg.bvh_leaf_max <= 0) ctx->cfg.bvh_leaf_max = 2;   // measured: 2 beats 3 on every BASELINE scene (fewer primitive tests per ray for a few more nodes)
    ctx->num_sms = prop.multiProcessorCount;
    const size_t n = (size_t)cfg->width * cfg->height;
    cudaError_t e2 = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking);
    ctx->stream = ctx->own_stream;
    if (e2 == cudaSuccess) e2 = cudaEventCreate(&ctx->ev0);
    if (e2 == cudaSuccess) e2 = cudaEventCreate(&ctx->ev1);
    if (e2 == cudaSuccess) e2 = cudaMalloc(&ctx->fb.rng, n * 6 * sizeof(uint32_t));
    if (e2 == cudaSuccess) e2 = cudaMalloc(&ctx->fb.accum, n * sizeof(float4));
    if (e2 == cudaSuccess) e2 = cudaMalloc(&ctx->fb.rgba8, n * sizeof(uint32_t));
    if (e2 == cudaSuccess) e2 = cudaMalloc(&ctx->fb.counters, 8 * sizeof(unsigned long long));
    if (e2 == cudaSuccess) e2 = cudaMemsetAsync(ctx->fb.accum, 0, n * sizeof(float4), ctx->stream);
    if (e2 == cudaSuccess) e2 = cudaMemsetAsync(ctx->fb.rgba8, 0, n * sizeof(uint32_t), ctx->stream);
    if (e2 == cudaSuccess && cfg->mode == CRT_MODE_WAVEFRONT) e2 = wavefront_create(&ctx->wf, cfg->width, cfg->height);   // AUTO: on first use
    if (e2 == cudaSuccess) e2 = cudaStreamSynchronize(ctx->stream);
    if (e2 != cudaSuccess) {
        int rc = fail_cuda(nullptr, e2, "crt_create: allocation");
        crt_destroy(ctx);
        return rc;
    }
    *out_ctx = ctx;
    return CRT_OK;
}

void crt_destroy(crt_ctx* ctx) {
    if (!ctx) return;
    if (ctx->worker.joinable()) ctx->worker.join();   // an asynchronous render still in flight finishes first
    cudaSetDevice(ctx->cfg.device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    crt_comm_release_(ctx);
    free_scene(ctx);
    wavefront_destroy(ctx->wf);
    cudaFree(ctx->fb.rng); cudaFree(ctx->fb.accum); cudaFree(ctx->fb.rgba8); cudaFree(ctx->fb.counters);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    (void)cudaGetLastError();
    delete ctx;
}

int crt_init_rng(crt_ctx* ctx, uint64_t seed, uint64_t subsequence_base) {
    NOT_BUSY(ctx);
    if (!ctx) return CRT_E_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    CU(launch_rng_init(ctx->fb.rng, ctx->cfg.width, ctx->cfg.height, seed, subsequence_base, ctx->num_sms, ctx->stream));
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaEventElapsedTime(&ctx->stats.rng_init_ms, ctx->ev0, ctx->ev1));
    ctx->rng_ready = true;
    return CRT_OK;
}

int crt_set_scene(crt_ctx* ctx, const crt_scene* sc) {
    NOT_BUSY(ctx);
    if (!ctx || !sc) return ctx ? fail(ctx, CRT_E_INVALID, "crt_set_scene: null scene") : CRT_E_INVALID;
    const uint32_t T = sc->num_triangles, S = sc->num_spheres;
    if (T + S == 0) return fail(ctx, CRT_E_INVALID, "crt_set_scene: empty scene");
    if (T && (!sc->positions || !sc->indices || !sc->tri_material || sc->num_vertices == 0))
        return fail(ctx, CRT_E_INVALID, "crt_set_scene: triangles without vertex/index/material arrays");
    if (S && (!sc->spheres || !sc->sphere_material)) return fail(ctx, CRT_E_INVALID, "crt_set_scene: spheres without arrays");
    if (sc->num_materials && !sc->materials) return fail(ctx, CRT_E_INVALID, "crt_set_scene: null materials");
    for (size_t i = 0; i < (size_t)3 * T; i++)
        if (sc->indices[i] >= sc->num_vertices) return fail(ctx, CRT_E_INVALID, "crt_set_scene: vertex index out of range");
    // Attenuations must be finite and >= 0: the path throughput is their product, and the wavefront kernels keep a
    // per-pixel flag in the sign bit of throughput.x (crt_wavefront.cu, PARKED_SIGN). A negative or NaN albedo has no
    // physical meaning; -0.0 is stored as +0.0 (same radiance bits: only products with 0 are affected).
    std::vector<MatRec> mats;
    try {
        mats.reserve(sc->num_materials);
        for (uint32_t i = 0; i < sc->num_materials; i++) {
            MatRec m = make_mat_record(sc->materials[i]);
            if (m.type == CRT_MAT_LAMBERTIAN || m.type == CRT_MAT_METAL) {
                float* c[3] = {&m.a.x, &m.a.y, &m.a.z};
                for (float* v : c) {
                    if (!(*v >= 0.f) || !std::isfinite(*v)) return fail(ctx, CRT_E_INVALID, "crt_set_scene: albedo components must be finite and >= 0");
                    *v += 0.f;   // -0 -> +0
                }
            }
            mats.push_back(m);
        }
    } catch (const std::bad_alloc&) {
        return fail(ctx, CRT_E_NOMEM, "crt_set_scene: out of host memory");
    }
    // reference-exact mode: replay the reference builder's partition on the host (crt_refexact.cpp): tie keys, and the
    // triangles its zero-thickness leaves hide, which become degenerate (v0, v0, v0) here: same id, never hit
    std::vector<uint32_t> order, rank, idx_seen;
    uint32_t unseen_count = 0;
    if (ctx->cfg.reference_exact && T) {
        try {
            order.resize(T); rank.resize(T);
            std::vector<uint8_t> unseen(T);
            const int rrc = crt_scene_reference_order(sc, order.data(), rank.data(), unseen.data());
            if (rrc != CRT_OK) return fail(ctx, rrc, "crt_set_scene: reference-order replay failed (mesh_tri_offset must tile the triangles in order)");
            for (uint32_t t = 0; t < T; t++) unseen_count += unseen[t];
            if (unseen_count) {
                idx_seen.assign(sc->indices, sc->indices + (size_t)3 * T);
                for (uint32_t t = 0; t < T; t++) if (unseen[t]) { idx_seen[3 * (size_t)t + 1] = idx_seen[3 * (size_t)t]; idx_seen[3 * (size_t)t + 2] = idx_seen[3 * (size_t)t]; }
            }
        } catch (const std::bad_alloc&) {
            return fail(ctx, CRT_E_NOMEM, "crt_set_scene: out of host memory");
        }
    }
    const uint32_t* host_indices = idx_seen.empty() ? sc->indices : idx_seen.data();
    CU(cudaSetDevice(ctx->cfg.device));
    CU(cudaStreamSynchronize(ctx->stream));
    free_scene(ctx);
    ctx->num_tris = T;
    ctx->unseen_tris = unseen_count;
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    if (T) {
        CU(cudaMalloc(&ctx->d_pos, (size_t)sc->num_vertices * 3 * sizeof(float)));
        CU(cudaMalloc(&ctx->d_idx, (size_t)T * 3 * sizeof(uint32_t)));
        CU(cudaMalloc(&ctx->d_trimat, (size_t)T * sizeof(int32_t)));
        CU(cudaMemcpyAsync(ctx->d_pos, sc->positions, (size_t)sc->num_vertices * 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->d_idx, host_indices, (size_t)T * 3 * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
        if (!rank.empty()) {
            CU(cudaMalloc(&ctx->d_trikey, (size_t)T * sizeof(uint32_t)));
            CU(cudaMalloc(&ctx->d_key2id, (size_t)T * sizeof(uint32_t)));
            CU(cudaMemcpyAsync(ctx->d_trikey, rank.data(), (size_t)T * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
            CU(cudaMemcpyAsync(ctx->d_key2id, order.data(), (size_t)T * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
        }
        CU(cudaMemcpyAsync(ctx->d_trimat, sc->tri_material, (size_t)T * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    }
    if (S) {
        CU(cudaMalloc(&ctx->d_sph, (size_t)S * 4 * sizeof(float)));
        CU(cudaMalloc(&ctx->d_sphmat, (size_t)S * sizeof(int32_t)));
        CU(cudaMemcpyAsync(ctx->d_sph, sc->spheres, (size_t)S * 4 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaMemcpyAsync(ctx->d_sphmat, sc->sphere_material, (size_t)S * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    }
    ctx->num_mats = sc->num_materials;
    CU(cudaMalloc(&ctx->d_mats, (mats.size() ? mats.size() : 1) * sizeof(MatRec)));
    if (!mats.empty()) CU(cudaMemcpyAsync(ctx->d_mats, mats.data(), mats.size() * sizeof(MatRec), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaEventElapsedTime(&ctx->stats.scene_upload_ms, ctx->ev0, ctx->ev1));

    SceneInput in;
    in.positions = ctx->d_pos; in.indices = ctx->d_idx; in.tri_material = ctx->d_trimat;
    in.spheres = ctx->d_sph; in.sphere_material = ctx->d_sphmat; in.num_triangles = T; in.num_spheres = S;
    in.tri_key = ctx->d_trikey;
    cudaError_t ce = cudaSuccess;
    int rc = build_bvh_device(in, (uint32_t)ctx->cfg.bvh_leaf_max, ctx->stream, ctx->num_sms, ctx->bvh, &ctx->stats.bvh_build_ms, &ce);
    if (rc != CRT_OK) {
        free_scene(ctx);
        if (ce != cudaSuccess) return fail_cuda(ctx, ce, "crt_set_scene: BVH build");
        return fail(ctx, rc, "crt_set_scene: BVH build failed");
    }
    if (2 * ctx->bvh.depth + 2 > CRT_STACK_ENTRIES) {
        free_scene(ctx);
        return fail(ctx, CRT_E_LIMIT, "crt_set_scene: BVH too deep for the traversal stack");
    }
    ctx->stats.bvh_nodes = ctx->bvh.num_nodes;
    ctx->stats.bvh_prims = ctx->bvh.num_prims;
    ctx->stats.bvh_depth = ctx->bvh.depth;
    ctx->stats.bvh_giants = ctx->bvh.giant_count;
    ctx->stats.unseen_triangles = ctx->unseen_tris;
    ctx->scene_ready = true;
    return CRT_OK;
}

int crt_set_camera(crt_ctx* ctx, const crt_camera* cam) {
    NOT_BUSY(ctx);
    if (!ctx || !cam) return ctx ? fail(ctx, CRT_E_INVALID, "crt_set_camera: null camera") : CRT_E_INVALID;
    ctx->cam = make_camera(*cam);
    ctx->cam_ready = true;
    return CRT_OK;
}

static int render_impl(crt_ctx* ctx, const crt_render_params* p, bool counted) {
    if (!ctx || !p) return ctx ? fail(ctx, CRT_E_INVALID, "crt_render: null params") : CRT_E_INVALID;
    if (!ctx->scene_ready) return fail(ctx, CRT_E_INVALID, "crt_render: no scene (call crt_set_scene)");
    if (!ctx->cam_ready) return fail(ctx, CRT_E_INVALID, "crt_render: no camera (call crt_set_camera)");
    if (!ctx->rng_ready) return fail(ctx, CRT_E_INVALID, "crt_render: RNG not initialised (call crt_init_rng)");
    if (p->spp <= 0 || p->spp >= (1 << 24) || p->max_bounces <= 0 || p->max_bounces > 255 || p->min_bounces < 1)
        return fail(ctx, CRT_E_INVALID, "crt_render: need 0 < spp < 2^24, 0 < max_bounces <= 255, min_bounces >= 1");
    CU(cudaSetDevice(ctx->cfg.device));
    RenderArgs a;
    a.scene = scene_view(ctx);
    a.cam = ctx->cam;
    a.rc.width = ctx->cfg.width; a.rc.height = ctx->cfg.height; a.rc.spp = p->spp;
    a.rc.max_bounces = p->max_bounces; a.rc.min_bounces = p->min_bounces; a.rc.rr_max_prob = p->rr_max_prob;
    a.fb = ctx->fb;
    a.flags = p->flags;
    bool mega = ctx->cfg.mode == CRT_MODE_MEGAKERNEL;
    if (ctx->cfg.mode == CRT_MODE_AUTO)
        mega = !((long long)ctx->cfg.width * ctx->cfg.height >= 1500000ll && ctx->bvh.num_prims >= 4096u);
    if (!mega && !ctx->wf) CU(wavefront_create(&ctx->wf, ctx->cfg.width, ctx->cfg.height));
    CU(cudaMemsetAsync(ctx->fb.counters, 0, 8 * sizeof(unsigned long long), ctx->stream));
    uint32_t launches = 0;
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    KernelTimes kt;
    bool incomplete = false;
    if (mega) CU(launch_render_mega(a, counted, ctx->num_sms, ctx->stream, &launches));
    else CU(launch_render_wavefront(ctx->wf, a, counted, ctx->num_sms, ctx->stream, &launches, &kt, &incomplete));
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    unsigned long long c[4] = {0, 0, 0, 0};
    CU(cudaMemcpyAsync(c, ctx->fb.counters, sizeof(c), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    CU(cudaEventElapsedTime(&ctx->stats.render_ms, ctx->ev0, ctx->ev1));
    ctx->stats.rays = c[0];
    ctx->stats.rays_total += c[0];
    ctx->stats.samples = (uint64_t)ctx->cfg.width * ctx->cfg.height * (uint64_t)p->spp;
    ctx->stats.nodes_fetched = c[2];
    ctx->stats.prims_tested = c[3];
    ctx->stats.kernel_launches = launches;
    if (mega) { kt.extend_ms = ctx->stats.render_ms; kt.extend_launches = 1; }
    ctx->stats.extend_ms = kt.extend_ms; ctx->stats.shade_ms = kt.shade_ms; ctx->stats.tail_ms = kt.tail_ms;
    ctx->stats.extend_launches = kt.extend_launches; ctx->stats.shade_launches = kt.shade_launches;
    ctx->stats.begin_ms = kt.begin_ms; ctx->stats.compact_ms = kt.compact_ms; ctx->stats.host_syncs = kt.host_syncs + 1u; ctx->stats.iterations = kt.iterations; ctx->stats.graph_launches = kt.graph_launches;
    if (incomplete) return fail(ctx, CRT_E_LIMIT, "crt_render: iteration bound reached with pixels still unfinished (frame incomplete)");
    return CRT_OK;
}

int crt_render(crt_ctx* ctx, const crt_render_params* p) {
    NOT_BUSY(ctx);
    if (ctx && p && (p->flags & CRT_RENDER_ASYNC)) {
        // the wavefront loop needs the host (it reads queue lengths back between batches), so "asynchronous" means a
        // host thread of the library's own drives the frame; the caller gets its thread back at once
        if (ctx->worker.joinable()) ctx->worker.join();
        ctx->async_params = *p;
        ctx->async_params.flags &= ~(uint32_t)CRT_RENDER_ASYNC;
        ctx->async_pending = true;
        try {
            ctx->worker = std::thread([ctx] { ctx->async_rc = render_impl(ctx, &ctx->async_params, false); });
        } catch (...) {
            ctx->async_pending = false;
            return fail(ctx, CRT_E_NOMEM, "crt_render: could not start the render thread");
        }
        return CRT_OK;
    }
    return render_impl(ctx, p, false);
}
int crt_render_counted(crt_ctx* ctx, const crt_render_params* p) { NOT_BUSY(ctx); return render_impl(ctx, p, true); }

int crt_wait(crt_ctx* ctx) {
    if (!ctx) return CRT_E_INVALID;
    if (!ctx->async_pending) return CRT_OK;
    if (ctx->worker.joinable()) ctx->worker.join();
    ctx->async_pending = false;
    return ctx->async_rc;
}

int crt_primary_ids(crt_ctx* ctx, uint64_t seed, uint64_t subsequence_base, int32_t* out_ids, float* out_t) {
    NOT_BUSY(ctx);
    if (!ctx || !out_ids) return ctx ? fail(ctx, CRT_E_INVALID, "crt_primary_ids: null output") : CRT_E_INVALID;
    if (!ctx->scene_ready || !ctx->cam_ready) return fail(ctx, CRT_E_INVALID, "crt_primary_ids: scene and camera must be set");
    CU(cudaSetDevice(ctx->cfg.device));
    const size_t n = (size_t)ctx->cfg.width * ctx->cfg.height;
    int32_t* d_ids = nullptr;
    float* d_t = nullptr;
    CU(cudaMalloc(&d_ids, n * sizeof(int32_t)));
    cudaError_t e = out_t ? cudaMalloc(&d_t, n * sizeof(float)) : cudaSuccess;
    if (e == cudaSuccess) e = launch_primary(scene_view(ctx), ctx->cam, ctx->cfg.width, ctx->cfg.height, seed, subsequence_base, d_ids, d_t, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_ids, d_ids, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && out_t) e = cudaMemcpyAsync(out_t, d_t, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_ids);
    cudaFree(d_t);
    if (e != cudaSuccess) return fail_cuda(ctx, e, "crt_primary_ids");
    return CRT_OK;
}

int crt_primary_ids_static(crt_ctx* ctx, uint64_t seed, uint64_t subsequence_base, int32_t* out_ids, float* out_t) {
    NOT_BUSY(ctx);
    if (!ctx || !out_ids) return ctx ? fail(ctx, CRT_E_INVALID, "crt_primary_ids_static: null output") : CRT_E_INVALID;
    if (!ctx->scene_ready || !ctx->cam_ready) return fail(ctx, CRT_E_INVALID, "crt_primary_ids_static: scene and camera must be set");
    const size_t n = (size_t)ctx->cfg.width * ctx->cfg.height;
    if (n > (size_t)CRT_PRIMARY_STATIC_MAX_PIXELS)
        return fail(ctx, CRT_E_LIMIT, "crt_primary_ids_static: " + std::to_string(n) + " pixels exceed the static scratch of " +
                                          std::to_string((size_t)CRT_PRIMARY_STATIC_MAX_PIXELS) + " pixels (CRT_PRIMARY_STATIC_MAX_PIXELS); use crt_primary_ids");
    CU(cudaSetDevice(ctx->cfg.device));
    int32_t* d_ids = nullptr;
    float* d_t = nullptr;
    cudaError_t e = launch_primary_static(scene_view(ctx), ctx->cam, ctx->cfg.width, ctx->cfg.height, seed, subsequence_base,
                                          out_t != nullptr, ctx->cfg.device, ctx->stream, &d_ids, &d_t);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_ids, d_ids, n * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && out_t) e = cudaMemcpyAsync(out_t, d_t, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return fail_cuda(ctx, e, "crt_primary_ids_static");
    return CRT_OK;
}

int crt_get_image_rgba8(crt_ctx* ctx, void* host, void** out_dev) {
    NOT_BUSY(ctx);
    if (!ctx) return CRT_E_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    if (out_dev) *out_dev = ctx->fb.rgba8;
    if (host) {
        CU(cudaMemcpyAsync(host, ctx->fb.rgba8, (size_t)ctx->cfg.width * ctx->cfg.height * 4, cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return CRT_OK;
}

int crt_get_accum(crt_ctx* ctx, float* host, void** out_dev) {
    NOT_BUSY(ctx);
    if (!ctx) return CRT_E_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    if (out_dev) *out_dev = ctx->fb.accum;
    if (host) {
        CU(cudaMemcpyAsync(host, ctx->fb.accum, (size_t)ctx->cfg.width * ctx->cfg.height * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return CRT_OK;
}

int crt_set_accum(crt_ctx* ctx, const float* host) {
    NOT_BUSY(ctx);
    if (!ctx || !host) return ctx ? fail(ctx, CRT_E_INVALID, "crt_set_accum: null buffer") : CRT_E_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    CU(cudaMemcpyAsync(ctx->fb.accum, host, (size_t)ctx->cfg.width * ctx->cfg.height * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}

int crt_tonemap(crt_ctx* ctx) {
    NOT_BUSY(ctx);
    if (!ctx) return CRT_E_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    CU(launch_tonemap(ctx->fb, ctx->cfg.width, ctx->cfg.height, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return CRT_OK;
}

int crt_get_rng_state(crt_ctx* ctx, uint32_t* host, void** out_dev) {
    NOT_BUSY(ctx);
    if (!ctx) return CRT_E_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    if (out_dev) *out_dev = ctx->fb.rng;
    if (host) {
        CU(cudaMemcpyAsync(host, ctx->fb.rng, (size_t)ctx->cfg.width * ctx->cfg.height * 6 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return CRT_OK;
}

int crt_set_rng_state(crt_ctx* ctx, const uint32_t* host) {
    NOT_BUSY(ctx);
    if (!ctx || !host) return ctx ? fail(ctx, CRT_E_INVALID, "crt_set_rng_state: null buffer") : CRT_E_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    CU(cudaMemcpyAsync(ctx->fb.rng, host, (size_t)ctx->cfg.width * ctx->cfg.height * 6 * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->rng_ready = true;
    return CRT_OK;
}

int crt_get_stats(crt_ctx* ctx, crt_stats* out) {
    if (!ctx || !out) return CRT_E_INVALID;
    *out = ctx->stats;
    return CRT_OK;
}

int crt_set_stream(crt_ctx* ctx, void* stream) {
    NOT_BUSY(ctx);
    if (!ctx) return CRT_E_INVALID;
    CU(cudaSetDevice(ctx->cfg.device));
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->stream = stream ? (cudaStream_t)stream : ctx->own_stream;
    return CRT_OK;
}

int crt_get_stream(crt_ctx* ctx, void** out_stream) {
    if (!ctx || !out_stream) return CRT_E_INVALID;
    *out_stream = (void*)ctx->stream;
    return CRT_OK;
}

}  // extern "C"
