// scene_gen.h — procedural stand-ins for the reference's (absent) assets.
//
// The reference loads assets/models/CornellBox-Original.obj and assets/models/bunny.obj
// (reference: CudaRayTracer/src/SceneManager.h:101-103); `assets/` is git-ignored upstream
// and absent, so the BASELINE.json configs are generated here, already in the
// reference's post-normalisation units (max extent 0.6 about the origin,
// reference: src/SceneManager.h:307-325).
//
// Plain host C++ (no CUDA): the same translation unit is linked into libcrt_b200.so,
// into the reference shim (oracle/ref_shim) and is what the tests hand to the oracle,
// so all three see bit-identical arrays.
#pragma once
#include <cstdint>
#include <string>
#include <vector>
#include "../../include/crt_api.h"

namespace crt {

struct SceneData {
    std::vector<float>        positions;      // xyz
    std::vector<uint32_t>     indices;        // 3 per triangle, global vertex numbers
    std::vector<int32_t>      tri_material;
    std::vector<float>        spheres;        // cx cy cz r
    std::vector<int32_t>      sphere_material;
    std::vector<crt_material> materials;
    std::vector<uint32_t>     mesh_tri_offset; // num_meshes+1

    uint32_t num_vertices() const { return (uint32_t)(positions.size() / 3); }
    uint32_t num_triangles() const { return (uint32_t)(indices.size() / 3); }
    uint32_t num_spheres() const { return (uint32_t)(spheres.size() / 4); }
    uint32_t num_meshes() const { return mesh_tri_offset.empty() ? 0u : (uint32_t)mesh_tri_offset.size() - 1u; }

    // Non-owning view (valid while *this is alive and unmodified).
    crt_scene view() const;
};

// names: "cornell" (C1), "cornell_blob" (C2/C5), "spheres" (C3), "terrain" (C4).
// param: 0 = BASELINE size. cornell_blob: param = latitude bands (default 136, longitude
// segments = ceil(bands*256/136)); terrain: param = grid cells along x (default 1582).
// Returns false for an unknown name.
bool make_builtin_scene(const std::string& name, int param, SceneData& out, crt_camera& cam);

// The two spheres createRandomWorld always appends (reference: src/CUDAKernels.h:69-73)
// together with their two materials; returns the index of the ground material.
int append_reference_spheres(SceneData& s);

// Camera helper: yaw/pitch (degrees) such that the reference camera, whose view direction is
// -front(yaw,pitch) (reference: src/Core/Camera.cuh:161-169,179), looks from `from` at `to`.
void look_at_yaw_pitch(const float from[3], const float to[3], float& yaw_deg, float& pitch_deg);

}  // namespace crt
