// crt_obj.cpp — OBJ/MTL ingest with the reference's scene-assembly rules.
//
// Replaces SceneManager::loadObject + initMeshes (reference: CudaRayTracer/src/SceneManager.h:100-329).
// The reference parses with tinyobjloader (un-vendored, v1.0.x API, SceneManager.h:215); this is an
// own parser of the subset that matters to the render path: v / f (any polygon, fan-triangulated;
// v, v/vt, v//vn, v/vt/vn; negative = relative indices) / usemtl / mtllib, and in MTL newmtl, Kd, Ks,
// Ke, d, Tr, Ni, Ns, Pr. Parsing parity with tinyobjloader is unpinned (no reference fixtures exist).
//
// Reference behaviour reproduced on purpose:
//   - one mesh per FILE (all shapes merged), vertex array sized 3x the vertex count with the unused
//     tail left at zero (SceneManager.h:253; the zeros take part in the normalisation below);
//   - material class: emission > 0 -> light; d < 1 -> dielectric(Ni); Ks.r > 0 -> metal with
//     roughness Pr, or sqrt(2/(Ns+2)) when Pr is 0; else Lambertian(Kd)            (:222-247)
//   - face material id = file-local id, out-of-range -> 0                               (:259-266)
//   - after EACH file, all meshes loaded so far are re-centred and re-scaled to a 0.6 max extent (:307-325)
//   - mesh i's material ids are offset by the number of DISTINCT ids used by mesh i-1     (:143-145,177)
//   - createRandomWorld's ground + metal spheres and their materials go last (CUDAKernels.h:69-73)
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <new>
#include <set>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/crt_api.h"
#include "scene_gen.h"

namespace crt {
crt_scene* scene_holder_release(SceneData&& data);
}

namespace {

struct MtlEntry {   // tinyobj material_t defaults: colours 0, dissolve 1, ior 1, shininess 1, roughness 0
    std::string name;
    float kd[3] = {0, 0, 0}, ks[3] = {0, 0, 0}, ke[3] = {0, 0, 0};
    float dissolve = 1.f, ior = 1.f, shininess = 1.f, roughness = 0.f;
};

bool parse_mtl(const std::string& path, std::vector<MtlEntry>& out, std::map<std::string, int>& by_name) {
    std::ifstream f(path);
    if (!f) return false;
    std::string line;
    MtlEntry* cur = nullptr;
    while (std::getline(f, line)) {
        std::istringstream ss(line);
        std::string key;
        if (!(ss >> key) || key[0] == '#') continue;
        if (key == "newmtl") {
            std::string name;
            ss >> name;
            by_name[name] = (int)out.size();
            out.emplace_back();
            out.back().name = name;
            cur = &out.back();
        } else if (cur) {
            if (key == "Kd") ss >> cur->kd[0] >> cur->kd[1] >> cur->kd[2];
            else if (key == "Ks") ss >> cur->ks[0] >> cur->ks[1] >> cur->ks[2];
            else if (key == "Ke") ss >> cur->ke[0] >> cur->ke[1] >> cur->ke[2];
            else if (key == "d") ss >> cur->dissolve;
            else if (key == "Tr") { float tr = 0.f; ss >> tr; cur->dissolve = 1.f - tr; }
            else if (key == "Ni") ss >> cur->ior;
            else if (key == "Ns") ss >> cur->shininess;
            else if (key == "Pr") ss >> cur->roughness;
        }
    }
    return true;
}

crt_material classify(const MtlEntry& m) {   // SceneManager.h:222-247
    crt_material r;
    std::memset(&r, 0, sizeof(r));
    int type = CRT_MAT_LAMBERTIAN;
    if (m.ke[0] > 0.f || m.ke[1] > 0.f || m.ke[2] > 0.f) type = CRT_MAT_DIFFUSE_LIGHT;
    else if (m.dissolve < 1.f) type = CRT_MAT_DIELECTRIC;
    else if (m.ks[0] > 0.f) type = CRT_MAT_METAL;
    float rough = 0.f;
    if (type == CRT_MAT_METAL) rough = m.roughness > 0.f ? m.roughness : sqrtf(2.f / (m.shininess + 2.f));
    r.type = type;
    for (int k = 0; k < 3; k++) { r.albedo[k] = m.kd[k]; r.emission[k] = m.ke[k]; }
    r.roughness = rough;
    r.ior = type == CRT_MAT_DIELECTRIC ? m.ior : 1.f;
    return r;
}

struct FileMesh {
    std::vector<float> verts;         // xyz, (3 * nverts) entries as in the reference
    std::vector<uint32_t> indices;
    std::vector<int> face_mat;        // file-local, clamped
};

int resolve_index(long idx, size_t count) {   // OBJ: 1-based, negative = relative to the end
    if (idx > 0) return (int)(idx - 1);
    if (idx < 0) return (int)((long)count + idx);
    return -1;
}

bool parse_obj(const std::string& path, FileMesh& mesh, std::vector<MtlEntry>& mtls, std::string& err) {
    std::ifstream f(path);
    if (!f) { err = "cannot open " + path; return false; }
    std::string base;
    size_t slash = path.find_last_of("/\\");
    if (slash != std::string::npos) base = path.substr(0, slash + 1);
    std::vector<float> pos;
    std::map<std::string, int> by_name;
    struct Face { int v[3]; int mat; };
    std::vector<Face> faces;
    int cur_mat = -1;
    std::string line;
    size_t lineno = 0;
    while (std::getline(f, line)) {
        lineno++;
        std::istringstream ss(line);
        std::string key;
        if (!(ss >> key) || key[0] == '#') continue;
        if (key == "v") {
            float x = 0, y = 0, z = 0;
            if (!(ss >> x >> y >> z)) { err = path + ":" + std::to_string(lineno) + ": bad vertex"; return false; }
            pos.push_back(x); pos.push_back(y); pos.push_back(z);
        } else if (key == "f") {
            std::vector<int> poly;
            std::string tok;
            while (ss >> tok) {
                long vi = strtol(tok.c_str(), nullptr, 10);   // the part before the first '/'
                int r = resolve_index(vi, pos.size() / 3);
                if (r < 0 || (size_t)r >= pos.size() / 3) { err = path + ":" + std::to_string(lineno) + ": vertex index out of range"; return false; }
                poly.push_back(r);
            }
            if (poly.size() < 3) { err = path + ":" + std::to_string(lineno) + ": face with fewer than 3 vertices"; return false; }
            for (size_t k = 2; k < poly.size(); k++) faces.push_back(Face{{poly[0], poly[k - 1], poly[k]}, cur_mat});
        } else if (key == "usemtl") {
            std::string name;
            ss >> name;
            auto it = by_name.find(name);
            cur_mat = it == by_name.end() ? -1 : it->second;
        } else if (key == "mtllib") {
            std::string name;
            ss >> name;
            parse_mtl(base + name, mtls, by_name);   // a missing MTL is a warning upstream, not an error
        }
    }
    const size_t nverts = pos.size() / 3;
    mesh.verts.assign(nverts * 3 * 3, 0.f);   // resize(attrib.vertices.size()) Vertex entries, zero-filled
    for (const Face& fc : faces) {
        for (int k = 0; k < 3; k++) {
            int vi = fc.v[k];
            mesh.indices.push_back((uint32_t)vi);
            mesh.verts[(size_t)vi * 3] = pos[(size_t)vi * 3];
            mesh.verts[(size_t)vi * 3 + 1] = pos[(size_t)vi * 3 + 1];
            mesh.verts[(size_t)vi * 3 + 2] = pos[(size_t)vi * 3 + 2];
        }
        mesh.face_mat.push_back(fc.mat);
    }
    return true;
}

}  // namespace

extern "C" int crt_scene_load_obj(const char* const* paths, int32_t num_paths, int32_t add_reference_spheres,
                                  crt_scene** out_scene, char* err_buf, size_t err_len) {
    auto set_err = [&](const std::string& m) { if (err_buf && err_len) { snprintf(err_buf, err_len, "%s", m.c_str()); } };
    if (!paths || num_paths <= 0 || !out_scene) { set_err("crt_scene_load_obj: bad arguments"); return CRT_E_INVALID; }
    *out_scene = nullptr;
    try {   // no exception crosses the C boundary
    std::vector<FileMesh> meshes;
    std::vector<crt_material> materials;
    for (int i = 0; i < num_paths; i++) {
        FileMesh m;
        std::vector<MtlEntry> mtls;
        std::string err;
        if (!paths[i] || !parse_obj(paths[i], m, mtls, err)) {
            set_err(err.empty() ? "crt_scene_load_obj: null path" : err);
            std::ifstream probe(paths[i] ? paths[i] : "");
            return probe ? CRT_E_PARSE : CRT_E_IO;
        }
        for (const MtlEntry& e : mtls) materials.push_back(classify(e));
        // face ids are checked against the scene-wide material count so far (SceneManager.h:262)
        for (int& id : m.face_mat) if (id < 0 || id >= (int)materials.size()) id = 0;
        meshes.push_back(std::move(m));
        // re-normalise every mesh loaded so far (SceneManager.h:307-325)
        float mn[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f}, mx[3] = {-3.402823466e+38f, -3.402823466e+38f, -3.402823466e+38f};
        for (const FileMesh& fm : meshes)
            for (size_t v = 0; v + 2 < fm.verts.size(); v += 3)
                for (int k = 0; k < 3; k++) { mn[k] = fminf(mn[k], fm.verts[v + k]); mx[k] = fmaxf(mx[k], fm.verts[v + k]); }
        float c[3], ext = 0.f;
        for (int k = 0; k < 3; k++) { c[k] = (mn[k] + mx[k]) * 0.5f; }
        ext = fmaxf(mx[0] - mn[0], fmaxf(mx[1] - mn[1], mx[2] - mn[2]));
        float scale = 0.6f / ext;
        for (FileMesh& fm : meshes)
            for (size_t v = 0; v + 2 < fm.verts.size(); v += 3)
                for (int k = 0; k < 3; k++) fm.verts[v + k] = (fm.verts[v + k] - c[k]) * scale;
    }
    crt::SceneData s;
    s.mesh_tri_offset.push_back(0u);
    size_t prev_distinct = 0;
    for (size_t i = 0; i < meshes.size(); i++) {
        const FileMesh& fm = meshes[i];
        const uint32_t vbase = s.num_vertices();
        const int offset = i == 0 ? 0 : (int)prev_distinct;
        s.positions.insert(s.positions.end(), fm.verts.begin(), fm.verts.end());
        for (uint32_t idx : fm.indices) s.indices.push_back(vbase + idx);
        for (int id : fm.face_mat) s.tri_material.push_back(id + offset);
        std::set<int> distinct(fm.face_mat.begin(), fm.face_mat.end());
        prev_distinct = distinct.size();
        s.mesh_tri_offset.push_back(s.num_triangles());
    }
    s.materials = materials;
    if (add_reference_spheres) crt::append_reference_spheres(s);
    crt_scene* v = crt::scene_holder_release(std::move(s));
    if (!v) { set_err("out of memory"); return CRT_E_NOMEM; }
    *out_scene = v;
    return CRT_OK;
    } catch (const std::bad_alloc&) {
        set_err("crt_scene_load_obj: out of host memory");
        return CRT_E_NOMEM;
    } catch (...) {
        set_err("crt_scene_load_obj: unexpected C++ exception");
        return CRT_E_INVALID;
    }
}
