// hostsim.cpp — sequential host instantiation of the product's CRT_HD code. TEST INFRASTRUCTURE.
//
// There is no GPU in the development container, so the BVH build functors, the traversal and the
// shading code (all written host/device in raytracer-cuda_b200/csrc/*.h) are instantiated here with
// a sequential execution policy and compared with the oracle by tests/test_hostsim.py. This is a
// debugging aid for the algorithms only: it is never linked into libcrt_b200.so, proves nothing
// about the CUDA build, and is not used by the GPU parity tests or by bench.py.
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <vector>

#include "../../raytracer-cuda_b200/csrc/crt_bvh_driver.h"
#include "../../raytracer-cuda_b200/csrc/crt_camera_host.h"
#include "../../raytracer-cuda_b200/csrc/crt_shade.h"

using namespace crt;

struct HostExec {
    template <class T> T* alloc(size_t n) { return (T*)malloc((n ? n : 1) * sizeof(T)); }
    void release(void* p) { free(p); }
    template <class F> void for_n(int n, const F& f) { for (int i = 0; i < n; i++) f(i); }
    void sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t* keys_tmp, uint32_t* vals_tmp, int n, uint64_t** ks, uint32_t** vs) {
        std::vector<uint32_t> order((size_t)n);
        std::iota(order.begin(), order.end(), 0u);
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
        for (int i = 0; i < n; i++) { keys_tmp[i] = keys[order[(size_t)i]]; vals_tmp[i] = vals[order[(size_t)i]]; }
        *ks = keys_tmp;
        *vs = vals_tmp;
    }
    void exclusive_scan(const uint32_t* in, uint32_t* out, int n) { uint32_t s = 0; for (int i = 0; i < n; i++) { out[i] = s; s += in[i]; } }
    void fill_u32(uint32_t* p, uint32_t v, size_t n) { for (size_t i = 0; i < n; i++) p[i] = v; }
    void upload(void* d, const void* s, size_t b) { memcpy(d, s, b); }
    void download(void* d, const void* s, size_t b) { memcpy(d, s, b); }
    void ploc_small(const PlocSmallArgs& a) { ploc_small_on_host(*this, a); }
    void reserve_tmp(size_t) {}
    template <class T> T* alloc_tmp(size_t n) { return alloc<T>(n); }
    void release_tmp(void* p) { release(p); }
};

struct hs_scene {
    WideBvh bvh;
    std::vector<MatRec> mats;
    SceneView view;
};

extern "C" {

hs_scene* hs_scene_create(const crt_scene* sc, int leaf_max) {
    HostExec ex;
    SceneInput in;
    in.positions = sc->positions; in.indices = sc->indices; in.tri_material = sc->tri_material;
    in.spheres = sc->spheres; in.sphere_material = sc->sphere_material;
    in.num_triangles = sc->num_triangles; in.num_spheres = sc->num_spheres;
    hs_scene* s = new hs_scene();
    if (build_wide_bvh(ex, in, (uint32_t)(leaf_max > 0 ? leaf_max : 3), s->bvh) != CRT_OK) { delete s; return nullptr; }
    for (uint32_t i = 0; i < sc->num_materials; i++) s->mats.push_back(make_mat_record(sc->materials[i]));
    s->view.nodes = s->bvh.nodes; s->view.prims = s->bvh.prims; s->view.mats = s->mats.data();
    s->view.num_mats = (uint32_t)s->mats.size(); s->view.num_nodes = s->bvh.num_nodes; s->view.num_prims = s->bvh.num_prims;
    s->view.giant_base = s->bvh.giant_base; s->view.giant_count = s->bvh.giant_count;
    s->view.key_to_id = nullptr; s->view.num_tris = sc->num_triangles;
    return s;
}

void hs_scene_destroy(hs_scene* s) { if (s) { free(s->bvh.nodes); free(s->bvh.prims); delete s; } }

void hs_scene_info(const hs_scene* s, uint32_t out[4]) {
    out[0] = s->bvh.num_nodes; out[1] = s->bvh.num_prims; out[2] = s->bvh.depth; out[3] = s->bvh.ploc_rounds;
}
uint32_t hs_scene_giants(const hs_scene* s) { return s->bvh.giant_count; }

const void* hs_scene_nodes(const hs_scene* s) { return s->bvh.nodes; }
const void* hs_scene_prims(const hs_scene* s) { return s->bvh.prims; }

// rng6: per pixel v0..v4,d (AoS, W*H*6), supplied by the caller (from the oracle's XORWOW init)
int hs_primary(const hs_scene* s, const crt_camera* cam, int W, int H, const uint32_t* rng6, int32_t* ids, float* t,
               uint64_t* nodes, uint64_t* prims) {
    Camera k = make_camera(*cam);
    TraceCounters ctr{0, 0};
    uint64_t tn = 0, tp = 0;
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            size_t pix = (size_t)y * W + x;
            Rng r{rng6[pix * 6], rng6[pix * 6 + 1], rng6[pix * 6 + 2], rng6[pix * 6 + 3], rng6[pix * 6 + 4], rng6[pix * 6 + 5]};
            vec3 o, d;
            camera_ray(k, x, y, W, H, r, o, d);
            HitRec h;
            ctr.nodes = ctr.prims = 0;
            bool hit = trace_closest<true>(s->view, o, d, 0.001f, INFINITY, h, &ctr);
            tn += ctr.nodes; tp += ctr.prims;
            ids[pix] = hit ? h.id : -1;
            t[pix] = hit ? h.t : INFINITY;
        }
    if (nodes) *nodes = tn;
    if (prims) *prims = tp;
    return 0;
}

int hs_render(const hs_scene* s, const crt_camera* cam, int W, int H, int spp, int max_bounces, uint32_t* rng6,
              float* radiance, uint8_t* rgba8, uint64_t* rays, uint64_t* nodes, uint64_t* prims) {
    Camera k = make_camera(*cam);
    RenderConst rc{W, H, spp, max_bounces, 3, 0.95f};
    const float scale = 1.f / (float)spp;
    uint64_t tr = 0, tn = 0, tp = 0;
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            size_t pix = (size_t)y * W + x;
            Rng r{rng6[pix * 6], rng6[pix * 6 + 1], rng6[pix * 6 + 2], rng6[pix * 6 + 3], rng6[pix * 6 + 4], rng6[pix * 6 + 5]};
            vec3 sum = mk3(0.f, 0.f, 0.f);
            TraceCounters ctr{0, 0};
            uint32_t nr = 0;
            for (int i = 0; i < spp; i++) {
                vec3 o, d;
                camera_ray(k, x, y, W, H, r, o, d);
                vec3 c = path_radiance<true>(s->view, rc, o, d, r, nr, &ctr);
                sum = v_add(sum, c);
            }
            tr += nr; tn += ctr.nodes; tp += ctr.prims;
            vec3 outc = mk3(f_mul(scale, sum.x), f_mul(scale, sum.y), f_mul(scale, sum.z));
            if (radiance) { radiance[3 * pix] = outc.x; radiance[3 * pix + 1] = outc.y; radiance[3 * pix + 2] = outc.z; }
            if (rgba8) { uint32_t p = pack_rgba8(outc); memcpy(rgba8 + 4 * pix, &p, 4); }
            uint32_t st[6] = {r.v0, r.v1, r.v2, r.v3, r.v4, r.d};
            memcpy(rng6 + pix * 6, st, sizeof(st));
        }
    if (rays) *rays = tr;
    if (nodes) *nodes = tn;
    if (prims) *prims = tp;
    return 0;
}

}  // extern "C"
