// crt_host.cpp — host-only parts of the C-ABI: built-in scenes, image files, checkpoints.
// (OBJ/MTL ingest lives in crt_obj.cpp.) Everything here goes through the public entry points
// of crt_api.cu; no device code.
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/crt_api.h"
#include "scene_gen.h"

extern "C" int crt_ctx_size_(const crt_ctx* ctx, int* w, int* h);   // crt_api.cu (not exported in the header)
extern "C" int crt_ctx_fail_(crt_ctx* ctx, int code, const char* msg);

namespace {
// The crt_scene view is the first member so crt_scene_free can recover the holder.
struct SceneHolder {
    crt_scene view;
    crt::SceneData data;
};
}  // namespace

namespace crt {
crt_scene* scene_holder_release(SceneData&& data) {
    SceneHolder* h = new (std::nothrow) SceneHolder();
    if (!h) return nullptr;
    h->data = std::move(data);
    h->view = h->data.view();
    return &h->view;
}
}  // namespace crt

// No exception crosses the C boundary: a std::bad_alloc from the host-side staging vectors (a 2560x1440 checkpoint
// stages ~150 MB) becomes CRT_E_NOMEM, anything else CRT_E_INVALID.
#define CRT_TRY try {
#define CRT_CATCH(ctx, what)                                                                            \
    } catch (const std::bad_alloc&) { return crt_ctx_fail_(ctx, CRT_E_NOMEM, what ": out of host memory");  \
    } catch (...) { return crt_ctx_fail_(ctx, CRT_E_INVALID, what ": unexpected C++ exception"); }

extern "C" {

int crt_scene_builtin(const char* name, int32_t param, crt_scene** out_scene, crt_camera* out_cam) {
    CRT_TRY
    if (!name || !out_scene) return CRT_E_INVALID;
    *out_scene = nullptr;
    crt::SceneData data;
    crt_camera cam;
    if (!crt::make_builtin_scene(name, param, data, cam)) return CRT_E_INVALID;
    crt_scene* v = crt::scene_holder_release(std::move(data));
    if (!v) return CRT_E_NOMEM;
    *out_scene = v;
    if (out_cam) *out_cam = cam;
    return CRT_OK;
    CRT_CATCH(nullptr, "crt_scene_builtin")
}

void crt_scene_free(crt_scene* scene) {
    if (!scene) return;
    delete reinterpret_cast<SceneHolder*>(scene);
}

// Replaces WindowManager::drawFrame's D2H copy + flipVertically (reference:
// CudaRayTracer/src/WindowManager.h:79-93): the file is written top row first.
int crt_write_ppm(crt_ctx* ctx, const char* path) {
    CRT_TRY
    int w = 0, h = 0;
    if (!ctx || !path || crt_ctx_size_(ctx, &w, &h) != CRT_OK) return CRT_E_INVALID;
    std::vector<unsigned char> img((size_t)w * h * 4);
    int rc = crt_get_image_rgba8(ctx, img.data(), nullptr);
    if (rc != CRT_OK) return rc;
    FILE* f = fopen(path, "wb");
    if (!f) return crt_ctx_fail_(ctx, CRT_E_IO, "crt_write_ppm: cannot open file");
    fprintf(f, "P6\n%d %d\n255\n", w, h);
    std::vector<unsigned char> row((size_t)w * 3);
    for (int y = h - 1; y >= 0; y--) {
        for (int x = 0; x < w; x++) memcpy(&row[(size_t)x * 3], &img[((size_t)y * w + x) * 4], 3);
        fwrite(row.data(), 1, row.size(), f);
    }
    bool ok = !ferror(f);
    fclose(f);
    return ok ? CRT_OK : crt_ctx_fail_(ctx, CRT_E_IO, "crt_write_ppm: write failed");
    CRT_CATCH(ctx, "crt_write_ppm")
}

int crt_write_pfm(crt_ctx* ctx, const char* path) {
    CRT_TRY
    int w = 0, h = 0;
    if (!ctx || !path || crt_ctx_size_(ctx, &w, &h) != CRT_OK) return CRT_E_INVALID;
    std::vector<float> acc((size_t)w * h * 4);
    int rc = crt_get_accum(ctx, acc.data(), nullptr);
    if (rc != CRT_OK) return rc;
    FILE* f = fopen(path, "wb");
    if (!f) return crt_ctx_fail_(ctx, CRT_E_IO, "crt_write_pfm: cannot open file");
    fprintf(f, "PF\n%d %d\n-1.0\n", w, h);   // negative scale = little endian; PFM rows are bottom-up
    std::vector<float> row((size_t)w * 3);
    for (int y = 0; y < h; y++) {
        for (int x = 0; x < w; x++) {
            const float* a = &acc[((size_t)y * w + x) * 4];
            float inv = a[3] > 0.f ? 1.f / a[3] : 0.f;
            row[(size_t)x * 3] = a[0] * inv; row[(size_t)x * 3 + 1] = a[1] * inv; row[(size_t)x * 3 + 2] = a[2] * inv;
        }
        fwrite(row.data(), sizeof(float), row.size(), f);
    }
    bool ok = !ferror(f);
    fclose(f);
    return ok ? CRT_OK : crt_ctx_fail_(ctx, CRT_E_IO, "crt_write_pfm: write failed");
    CRT_CATCH(ctx, "crt_write_pfm")
}

// Checkpoint file: "CRTCKPT1", width, height (int32), accumulation buffer (w*h*4 float),
// RNG planes (6*w*h uint32). The reference persists nothing.
int crt_checkpoint_save(crt_ctx* ctx, const char* path) {
    CRT_TRY
    int w = 0, h = 0;
    if (!ctx || !path || crt_ctx_size_(ctx, &w, &h) != CRT_OK) return CRT_E_INVALID;
    const size_t n = (size_t)w * h;
    std::vector<float> acc(n * 4);
    std::vector<uint32_t> rng(n * 6);
    int rc = crt_get_accum(ctx, acc.data(), nullptr);
    if (rc == CRT_OK) rc = crt_get_rng_state(ctx, rng.data(), nullptr);
    if (rc != CRT_OK) return rc;
    FILE* f = fopen(path, "wb");
    if (!f) return crt_ctx_fail_(ctx, CRT_E_IO, "crt_checkpoint_save: cannot open file");
    int32_t dims[2] = {w, h};
    fwrite("CRTCKPT1", 1, 8, f);
    fwrite(dims, sizeof(int32_t), 2, f);
    fwrite(acc.data(), sizeof(float), acc.size(), f);
    fwrite(rng.data(), sizeof(uint32_t), rng.size(), f);
    bool ok = !ferror(f);
    fclose(f);
    return ok ? CRT_OK : crt_ctx_fail_(ctx, CRT_E_IO, "crt_checkpoint_save: write failed");
    CRT_CATCH(ctx, "crt_checkpoint_save")
}

int crt_checkpoint_load(crt_ctx* ctx, const char* path) {
    CRT_TRY
    int w = 0, h = 0;
    if (!ctx || !path || crt_ctx_size_(ctx, &w, &h) != CRT_OK) return CRT_E_INVALID;
    FILE* f = fopen(path, "rb");
    if (!f) return crt_ctx_fail_(ctx, CRT_E_IO, "crt_checkpoint_load: cannot open file");
    char magic[8];
    int32_t dims[2] = {0, 0};
    const size_t n = (size_t)w * h;
    std::vector<float> acc(n * 4);
    std::vector<uint32_t> rng(n * 6);
    bool ok = fread(magic, 1, 8, f) == 8 && memcmp(magic, "CRTCKPT1", 8) == 0 && fread(dims, sizeof(int32_t), 2, f) == 2;
    if (ok && (dims[0] != w || dims[1] != h)) { fclose(f); return crt_ctx_fail_(ctx, CRT_E_INVALID, "crt_checkpoint_load: image size differs from the context"); }
    ok = ok && fread(acc.data(), sizeof(float), acc.size(), f) == acc.size() && fread(rng.data(), sizeof(uint32_t), rng.size(), f) == rng.size();
    fclose(f);
    if (!ok) return crt_ctx_fail_(ctx, CRT_E_IO, "crt_checkpoint_load: truncated or foreign file");
    int rc = crt_set_accum(ctx, acc.data());
    if (rc == CRT_OK) rc = crt_set_rng_state(ctx, rng.data());
    return rc;
    CRT_CATCH(ctx, "crt_checkpoint_load")
}

}  // extern "C"
