// crt_refexact.cpp — reference-exact mode: what the reference's OWN acceleration structure makes of a scene.
//
// Two behaviours of the reference depend on how its single-thread builder happens to partition the triangles of a mesh
// (reference: CudaRayTracer/src/Core/Mesh.cuh:121-219), not on the geometry alone:
//   * equal-t ties. Mesh::hit walks its BVH left first and accepts t <= closest (Mesh.cuh:55-110,293), so among hits
//     with bit-equal t the triangle that comes LAST in the builder's permuted index array wins; across objects the
//     later object wins (BVHNode::hit, Core/BVHNode.cuh:115-156: objects in insertion order, meshes then spheres).
//   * unseen triangles. A leaf's box is the plain min/max of its triangles (computeTrianglesAABB, Mesh.cuh:258-264:
//     no padToMinimums, Core/AABB.cuh:181-186), and AABB::hit rejects tmax <= tmin (AABB.cuh:135-136): a leaf whose
//     triangles all lie in one axis-aligned plane has a box of zero thickness that no ray ever enters. The reference
//     never sees those triangles (axis-aligned quads such as Cornell-box walls, when a leaf holds nothing else).
// This library traverses a different BVH, so by default it keeps the larger primitive id on a tie and sees every
// triangle. With crt_config.reference_exact it replays the reference's partition here, on the host (sequential like
// the original: ~5 s per million triangles), and
//   * uses each triangle's position in the reference's permuted order as the tie key (crt_trace.h test_prim),
//   * turns the unseen triangles into degenerate ones (they keep their ids; Moeller-Trumbore rejects them),
// so that primary ids and radiance equal the reference's on such scenes as well. Only the PARTITION is replayed: the
// arithmetic that decides it (centroids, the count x area cost with the compiled kernel's FMA contraction, the
// strict < comparisons) is restated operation by operation; tests pin it against the builder permutations dumped
// from the unmodified reference kernels (tests/golden/*_perm_i32.bin) through the oracle.
#include <cmath>
#include <cstdint>
#include <new>
#include <vector>

#include "../../include/crt_api.h"

namespace {

struct TriBox {
    float lo[3], hi[3];
};

struct Span {
    int start, end, node;
};

// one mesh: triangles [t0, t1) of the scene; order[] holds triangle ids and is permuted in place like m_Indices
void replay_mesh(const crt_scene& sc, uint32_t t0, uint32_t t1, uint32_t* order, uint8_t* unseen) {
    const int n = (int)(t1 - t0);
    if (n <= 0) return;
    const float third = 1.f / 3.f;
    std::vector<float> cen((size_t)n * 3);    // triangleCentroid, Mesh.cuh:251-256: ((p0 + p1) + p2) * (1/3)
    std::vector<TriBox> box((size_t)n);       // min / max of the three vertices (what expandTriangleAABB adds, :242-249)
    for (int k = 0; k < n; k++) {
        const uint32_t* ix = sc.indices + 3 * (size_t)(t0 + (uint32_t)k);
        const float* p0 = sc.positions + 3 * (size_t)ix[0];
        const float* p1 = sc.positions + 3 * (size_t)ix[1];
        const float* p2 = sc.positions + 3 * (size_t)ix[2];
        for (int a = 0; a < 3; a++) {
            cen[(size_t)k * 3 + a] = ((p0[a] + p1[a]) + p2[a]) * third;
            box[(size_t)k].lo[a] = fminf(fminf(p0[a], p1[a]), p2[a]);
            box[(size_t)k].hi[a] = fmaxf(fmaxf(p0[a], p1[a]), p2[a]);
        }
    }
    std::vector<int> perm((size_t)n);         // local triangle numbers in the builder's current order
    for (int k = 0; k < n; k++) perm[(size_t)k] = k;
    auto area2 = [](const float* lo, const float* hi) {   // AABB::area as compiled: 2 * fma(ez, ex, fma(ex, ey, ey * ez))
        const float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        return 2.0f * fmaf(ez, ex, fmaf(ex, ey, ey * ez));
    };
    const int max_nodes = 2 * n - 1;
    int next_node = 1;
    std::vector<Span> todo;
    todo.push_back({0, n, 0});
    while (!todo.empty()) {
        const Span cur = todo.back();
        todo.pop_back();
        const int count = cur.end - cur.start;
        if (count <= 10 || (next_node + 1) >= max_nodes) {   // span <= 30 indices, or the node budget (Mesh.cuh:148)
            // leaf: unseen when its exact box has no thickness along some axis
            float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
            for (int i = cur.start; i < cur.end; i++)
                for (int a = 0; a < 3; a++) {
                    lo[a] = fminf(lo[a], box[(size_t)perm[(size_t)i]].lo[a]);
                    hi[a] = fmaxf(hi[a], box[(size_t)perm[(size_t)i]].hi[a]);
                }
            const bool flat = lo[0] == hi[0] || lo[1] == hi[1] || lo[2] == hi[2];
            if (flat) for (int i = cur.start; i < cur.end; i++) unseen[t0 + (uint32_t)perm[(size_t)i]] = 1;
            continue;
        }
        int best_axis = 0;
        float best_pos = 0.f, best_cost = 1e30f;
        for (int axis = 0; axis < 3; axis++) {
            float mn = 1e30f, mx = -1e30f;
            for (int i = cur.start; i < cur.end; i++) {
                const float c = cen[(size_t)perm[(size_t)i] * 3 + axis];
                if (c < mn) mn = c;
                if (c > mx) mx = c;
            }
            const float pos = 0.5f * (mn + mx);
            // evaluateSAH, Mesh.cuh:222-240
            float llo[3] = {INFINITY, INFINITY, INFINITY}, lhi[3] = {-INFINITY, -INFINITY, -INFINITY};
            float rlo[3] = {INFINITY, INFINITY, INFINITY}, rhi[3] = {-INFINITY, -INFINITY, -INFINITY};
            int lc = 0, rc = 0;
            for (int i = cur.start; i < cur.end; i++) {
                const int t = perm[(size_t)i];
                const TriBox& b = box[(size_t)t];
                if (cen[(size_t)t * 3 + axis] < pos) {
                    lc++;
                    for (int a = 0; a < 3; a++) { llo[a] = fminf(llo[a], b.lo[a]); lhi[a] = fmaxf(lhi[a], b.hi[a]); }
                } else {
                    rc++;
                    for (int a = 0; a < 3; a++) { rlo[a] = fminf(rlo[a], b.lo[a]); rhi[a] = fmaxf(rhi[a], b.hi[a]); }
                }
            }
            float cost = fmaf((float)rc, area2(rlo, rhi), (float)lc * area2(llo, lhi));   // the kernel's FMUL + FFMA
            cost = cost < 1e-8f ? 1e-8f : cost;
            if (cost < best_cost) { best_cost = cost; best_axis = axis; best_pos = pos; }
        }
        int mid = cur.start;   // the reference's partition: a triangle left of the plane swaps with position `mid`
        for (int i = cur.start; i < cur.end; i++) {
            if (cen[(size_t)perm[(size_t)i] * 3 + best_axis] < best_pos) {
                const int tmp = perm[(size_t)mid]; perm[(size_t)mid] = perm[(size_t)i]; perm[(size_t)i] = tmp;
                mid++;
            }
        }
        const int left = next_node++, right = next_node++;
        todo.push_back({mid, cur.end, right});      // the left half is built first, like the reference's stack
        todo.push_back({cur.start, mid, left});
    }
    for (int k = 0; k < n; k++) order[t0 + (uint32_t)k] = t0 + (uint32_t)perm[(size_t)k];
}

}  // namespace

extern "C" int crt_scene_reference_order(const crt_scene* scene, uint32_t* out_order, uint32_t* out_rank, uint8_t* out_unseen) {
    if (!scene) return CRT_E_INVALID;
    const uint32_t T = scene->num_triangles;
    if (T && (!scene->positions || !scene->indices)) return CRT_E_INVALID;
    for (size_t i = 0; i < (size_t)3 * T; i++) if (scene->indices[i] >= scene->num_vertices) return CRT_E_INVALID;
    try {
        std::vector<uint32_t> order(T);
        std::vector<uint8_t> unseen(T, 0);
        if (scene->mesh_tri_offset && scene->num_meshes > 0) {
            uint32_t covered = 0;
            for (uint32_t m = 0; m < scene->num_meshes; m++) {
                const uint32_t a = scene->mesh_tri_offset[m], b = scene->mesh_tri_offset[m + 1];
                if (a != covered || b < a || b > T) return CRT_E_INVALID;   // the meshes tile [0, T) in order
                replay_mesh(*scene, a, b, order.data(), unseen.data());
                covered = b;
            }
            if (covered != T) return CRT_E_INVALID;
        } else {
            replay_mesh(*scene, 0, T, order.data(), unseen.data());
        }
        for (uint32_t k = 0; k < T; k++) {
            if (out_order) out_order[k] = order[k];
            if (out_rank) out_rank[order[k]] = k;
            if (out_unseen) out_unseen[k] = unseen[k];
        }
    } catch (const std::bad_alloc&) {
        return CRT_E_NOMEM;
    }
    return CRT_OK;
}
