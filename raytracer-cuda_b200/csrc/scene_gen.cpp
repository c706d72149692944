// scene_gen.cpp — see scene_gen.h. Host-only, deterministic, libm use limited to
// sinf/cosf/sqrtf/atan2f/asinf on the generating host (all consumers link this same file).
#include "scene_gen.h"

#include <cmath>
#include <cstring>
#include <random>

namespace crt {

crt_scene SceneData::view() const {
    crt_scene v;
    std::memset(&v, 0, sizeof(v));
    v.positions = positions.data();
    v.num_vertices = num_vertices();
    v.indices = indices.data();
    v.tri_material = tri_material.data();
    v.num_triangles = num_triangles();
    v.spheres = spheres.data();
    v.sphere_material = sphere_material.data();
    v.num_spheres = num_spheres();
    v.materials = materials.data();
    v.num_materials = (uint32_t)materials.size();
    v.mesh_tri_offset = mesh_tri_offset.empty() ? nullptr : mesh_tri_offset.data();
    v.num_meshes = num_meshes();
    return v;
}

namespace {

const float kPi = 3.14159265358979323846f;

crt_material make_mat(int type, float r, float g, float b, float rough = 0.f, float ior = 1.f) {
    crt_material m;
    std::memset(&m, 0, sizeof(m));
    m.type = type;
    if (type == CRT_MAT_DIFFUSE_LIGHT) {
        m.emission[0] = r; m.emission[1] = g; m.emission[2] = b;
    } else {
        m.albedo[0] = r; m.albedo[1] = g; m.albedo[2] = b;
    }
    m.roughness = rough;
    m.ior = ior;
    return m;
}

uint32_t add_vertex(SceneData& s, float x, float y, float z) {
    s.positions.push_back(x);
    s.positions.push_back(y);
    s.positions.push_back(z);
    return s.num_vertices() - 1;
}

void add_tri(SceneData& s, uint32_t a, uint32_t b, uint32_t c, int mat) {
    s.indices.push_back(a);
    s.indices.push_back(b);
    s.indices.push_back(c);
    s.tri_material.push_back(mat);
}

// quad a-b-c-d -> (a,b,c) + (a,c,d), four private vertices (like an OBJ face list
// triangulated as a fan, which is what tinyobjloader hands the reference).
void add_quad(SceneData& s, const float a[3], const float b[3], const float c[3], const float d[3], int mat) {
    uint32_t ia = add_vertex(s, a[0], a[1], a[2]);
    uint32_t ib = add_vertex(s, b[0], b[1], b[2]);
    uint32_t ic = add_vertex(s, c[0], c[1], c[2]);
    uint32_t id = add_vertex(s, d[0], d[1], d[2]);
    add_tri(s, ia, ib, ic, mat);
    add_tri(s, ia, ic, id, mat);
}

// Axis-aligned box rotated about y by `deg`, sitting at (cx, y0..y0+h, cz); 5 faces (no bottom).
void add_block(SceneData& s, float cx, float cz, float y0, float hx, float h, float hz, float deg, int mat) {
    float cs = cosf(deg * kPi / 180.f), sn = sinf(deg * kPi / 180.f);
    float px[4], pz[4];
    const float lx[4] = {-hx, hx, hx, -hx};
    const float lz[4] = {hz, hz, -hz, -hz};
    for (int i = 0; i < 4; i++) {
        px[i] = cx + cs * lx[i] + sn * lz[i];
        pz[i] = cz - sn * lx[i] + cs * lz[i];
    }
    float y1 = y0 + h;
    // top
    {
        float a[3] = {px[0], y1, pz[0]}, b[3] = {px[1], y1, pz[1]}, c[3] = {px[2], y1, pz[2]}, d[3] = {px[3], y1, pz[3]};
        add_quad(s, a, b, c, d, mat);
    }
    for (int i = 0; i < 4; i++) {
        int j = (i + 1) & 3;
        float a[3] = {px[i], y0, pz[i]}, b[3] = {px[j], y0, pz[j]}, c[3] = {px[j], y1, pz[j]}, d[3] = {px[i], y1, pz[i]};
        add_quad(s, a, b, c, d, mat);
    }
}

enum { MAT_WHITE = 0, MAT_RED = 1, MAT_GREEN = 2, MAT_LIGHT = 3 };

void cornell_materials(SceneData& s) {
    s.materials.push_back(make_mat(CRT_MAT_LAMBERTIAN, 0.725f, 0.71f, 0.68f));
    s.materials.push_back(make_mat(CRT_MAT_LAMBERTIAN, 0.63f, 0.065f, 0.05f));
    s.materials.push_back(make_mat(CRT_MAT_LAMBERTIAN, 0.14f, 0.45f, 0.091f));
    s.materials.push_back(make_mat(CRT_MAT_DIFFUSE_LIGHT, 17.f, 12.f, 4.f));
}

// Interior [-0.3,0.3]^3, open towards +z. 6 quads = 12 triangles.
void cornell_walls(SceneData& s) {
    const float L = -0.3f, H = 0.3f;
    {   // floor
        float a[3] = {L, L, H}, b[3] = {H, L, H}, c[3] = {H, L, L}, d[3] = {L, L, L};
        add_quad(s, a, b, c, d, MAT_WHITE);
    }
    {   // ceiling
        float a[3] = {L, H, L}, b[3] = {H, H, L}, c[3] = {H, H, H}, d[3] = {L, H, H};
        add_quad(s, a, b, c, d, MAT_WHITE);
    }
    {   // back wall
        float a[3] = {L, L, L}, b[3] = {H, L, L}, c[3] = {H, H, L}, d[3] = {L, H, L};
        add_quad(s, a, b, c, d, MAT_WHITE);
    }
    {   // left wall (red)
        float a[3] = {L, L, H}, b[3] = {L, L, L}, c[3] = {L, H, L}, d[3] = {L, H, H};
        add_quad(s, a, b, c, d, MAT_RED);
    }
    {   // right wall (green)
        float a[3] = {H, L, L}, b[3] = {H, L, H}, c[3] = {H, H, H}, d[3] = {H, H, L};
        add_quad(s, a, b, c, d, MAT_GREEN);
    }
    {   // ceiling light, 2 mm below the ceiling
        const float y = 0.298f;
        float a[3] = {-0.07f, y, -0.0575f}, b[3] = {0.07f, y, -0.0575f}, c[3] = {0.07f, y, 0.0575f}, d[3] = {-0.07f, y, 0.0575f};
        add_quad(s, a, b, c, d, MAT_LIGHT);
    }
}

void cornell_blocks(SceneData& s) {
    add_block(s, 0.10f, 0.10f, -0.3f, 0.09f, 0.18f, 0.09f, -17.f, MAT_WHITE);   // short block
    add_block(s, -0.10f, -0.09f, -0.3f, 0.09f, 0.36f, 0.09f, 18.f, MAT_WHITE);  // tall block
}

// ---- seeded value noise on a 3-D integer lattice -------------------------
inline uint32_t hash3(int x, int y, int z, uint32_t seed) {
    uint32_t h = seed;
    h ^= (uint32_t)x * 0x9E3779B1u; h = (h << 13) | (h >> 19); h *= 0x85EBCA6Bu;
    h ^= (uint32_t)y * 0xC2B2AE35u; h = (h << 15) | (h >> 17); h *= 0x27D4EB2Fu;
    h ^= (uint32_t)z * 0x165667B1u; h = (h << 11) | (h >> 21); h *= 0x9E3779B1u;
    h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
    return h;
}
inline float lattice(int x, int y, int z, uint32_t seed) {
    return (float)(hash3(x, y, z, seed) >> 8) * (1.0f / 16777216.0f) * 2.f - 1.f;   // [-1,1)
}
inline float smooth(float t) { return t * t * (3.f - 2.f * t); }
float value_noise(float x, float y, float z, uint32_t seed) {
    float fx = floorf(x), fy = floorf(y), fz = floorf(z);
    int ix = (int)fx, iy = (int)fy, iz = (int)fz;
    float tx = smooth(x - fx), ty = smooth(y - fy), tz = smooth(z - fz);
    float c[2][2][2];
    for (int a = 0; a < 2; a++)
        for (int b = 0; b < 2; b++)
            for (int d = 0; d < 2; d++) c[a][b][d] = lattice(ix + a, iy + b, iz + d, seed);
    float x00 = c[0][0][0] + tx * (c[1][0][0] - c[0][0][0]);
    float x10 = c[0][1][0] + tx * (c[1][1][0] - c[0][1][0]);
    float x01 = c[0][0][1] + tx * (c[1][0][1] - c[0][0][1]);
    float x11 = c[0][1][1] + tx * (c[1][1][1] - c[0][1][1]);
    float y0 = x00 + ty * (x10 - x00);
    float y1 = x01 + ty * (x11 - x01);
    return y0 + tz * (y1 - y0);
}
float fbm(float x, float y, float z, uint32_t seed, int octaves) {
    float amp = 0.5f, sum = 0.f, f = 1.f;
    for (int o = 0; o < octaves; o++) {
        sum += amp * value_noise(x * f, y * f, z * f, seed + (uint32_t)o * 101u);
        amp *= 0.5f;
        f *= 2.03f;
    }
    return sum;   // roughly [-1,1]
}

// Bunny stand-in: `bands` x `segs` parametric sphere, radius displaced by fBm, 2 triangles
// per cell -> bands*segs*2 triangles (136 x 256 x 2 = 69 632 ~ bunny.obj's 69 451).
// Latitude stops 0.02 rad short of the poles so no triangle is degenerate.
void add_blob(SceneData& s, float cx, float cy, float cz, float radius, int bands, int segs, uint32_t seed, int mat) {
    uint32_t base = s.num_vertices();
    const float t0 = 0.02f, t1 = kPi - 0.02f;
    for (int i = 0; i <= bands; i++) {
        float th = t0 + (t1 - t0) * (float)i / (float)bands;
        float st = sinf(th), ct = cosf(th);
        for (int j = 0; j < segs; j++) {
            float ph = 2.f * kPi * (float)j / (float)segs;
            float dx = st * cosf(ph), dy = ct, dz = st * sinf(ph);
            float r = radius * (1.f + 0.22f * fbm(dx * 2.3f + 11.f, dy * 2.3f + 5.f, dz * 2.3f + 7.f, seed, 4));
            add_vertex(s, cx + r * dx, cy + r * dy, cz + r * dz);
        }
    }
    for (int i = 0; i < bands; i++) {
        for (int j = 0; j < segs; j++) {
            int j1 = (j + 1) % segs;
            uint32_t a = base + (uint32_t)(i * segs + j), b = base + (uint32_t)(i * segs + j1);
            uint32_t c = base + (uint32_t)((i + 1) * segs + j1), d = base + (uint32_t)((i + 1) * segs + j);
            add_tri(s, a, b, c, mat);
            add_tri(s, a, c, d, mat);
        }
    }
}

// C4: height field over the Cornell floor, nx x nz cells, 2 triangles per cell.
void add_terrain(SceneData& s, int nx, int nz, uint32_t seed, int mat) {
    uint32_t base = s.num_vertices();
    const float L = -0.3f, W = 0.6f;
    for (int k = 0; k <= nz; k++) {
        float z = L + W * (float)k / (float)nz;
        for (int i = 0; i <= nx; i++) {
            float x = L + W * (float)i / (float)nx;
            float h = fbm(x * 9.f + 3.f, 0.5f, z * 9.f + 1.f, seed, 6);
            float ridge = fbm(x * 31.f, 7.5f, z * 31.f, seed + 977u, 3);
            float y = -0.3f + 0.02f + 0.16f * (0.5f + 0.5f * h) + 0.015f * ridge;
            add_vertex(s, x, y, z);
        }
    }
    uint32_t stride = (uint32_t)nx + 1u;
    for (int k = 0; k < nz; k++) {
        for (int i = 0; i < nx; i++) {
            uint32_t a = base + (uint32_t)k * stride + (uint32_t)i, b = a + 1, d = a + stride, c = d + 1;
            add_tri(s, a, d, c, mat);
            add_tri(s, a, c, b, mat);
        }
    }
}

void default_camera(crt_camera& cam) {
    std::memset(&cam, 0, sizeof(cam));
    cam.world_up[1] = 1.f;
    cam.yaw_deg = -90.f;
    cam.pitch_deg = 0.f;
    cam.aperture = 0.000001f;   // reference: src/EntryPoint.cu:20
}

}  // namespace

int append_reference_spheres(SceneData& s) {
    int ground = (int)s.materials.size();
    s.materials.push_back(make_mat(CRT_MAT_LAMBERTIAN, 0.5f, 0.5f, 0.5f));
    s.spheres.insert(s.spheres.end(), {0.f, -1000.f, 0.f, 999.f});
    s.sphere_material.push_back(ground);
    int metal = (int)s.materials.size();
    s.materials.push_back(make_mat(CRT_MAT_METAL, 0.7f, 0.6f, 0.5f, 0.0f));
    s.spheres.insert(s.spheres.end(), {0.2f, 0.2f, 0.f, 0.05f});
    s.sphere_material.push_back(metal);
    return ground;
}

void look_at_yaw_pitch(const float from[3], const float to[3], float& yaw_deg, float& pitch_deg) {
    float vx = to[0] - from[0], vy = to[1] - from[1], vz = to[2] - from[2];
    float len = sqrtf(vx * vx + vy * vy + vz * vz);
    vx /= len; vy /= len; vz /= len;
    // view = -front = (cos(yaw)cos(pitch), sin(pitch), sin(yaw)cos(pitch))
    pitch_deg = asinf(vy) * 180.f / kPi;
    yaw_deg = atan2f(vz, vx) * 180.f / kPi;
}

bool make_builtin_scene(const std::string& name, int param, SceneData& s, crt_camera& cam) {
    s = SceneData();
    default_camera(cam);
    if (name == "cornell") {
        // C1: walls + light + two blocks = 32 triangles, + the reference's two spheres.
        cornell_materials(s);
        cornell_walls(s);
        cornell_blocks(s);
        s.mesh_tri_offset = {0u, s.num_triangles()};
        append_reference_spheres(s);
        cam.position[0] = 0.f; cam.position[1] = 0.f; cam.position[2] = 0.95f;
        cam.vfov_deg = 40.f; cam.aspect = 1.f; cam.focus_dist = 0.95f;
        return true;
    }
    if (name == "cornell_blob") {
        // C2 / C5: C1 + bunny-sized blob standing on the tall block, grey Lambertian(0.5)
        // (the material an MTL-less bunny.obj lands on upstream).
        cornell_materials(s);
        cornell_walls(s);
        cornell_blocks(s);
        uint32_t t0 = s.num_triangles();
        int ground = append_reference_spheres(s);
        int bands = param > 0 ? param : 136;
        int segs = (bands * 256 + 135) / 136;
        add_blob(s, -0.10f, 0.06f + 0.095f, -0.09f, 0.085f, bands, segs, 42u, ground);
        s.mesh_tri_offset = {0u, t0, s.num_triangles()};
        // camera just inside the open side, every pixel sees the interior (as in the
        // upstream screenshot); vfov 80 as in src/EntryPoint.cu:19.
        cam.position[0] = 0.f; cam.position[1] = 0.f; cam.position[2] = 0.29f;
        cam.vfov_deg = 80.f; cam.aspect = 16.f / 9.f; cam.focus_dist = 0.45f;
        return true;
    }
    if (name == "spheres") {
        // C3: "Ray Tracing in One Weekend" final scene, host mt19937(1234).
        std::mt19937 rng(1234u);
        auto rnd = [&rng]() { return (float)(rng() >> 8) * (1.0f / 16777216.0f); };
        auto add_sphere = [&s](float x, float y, float z, float r, const crt_material& m) {
            s.spheres.insert(s.spheres.end(), {x, y, z, r});
            s.sphere_material.push_back((int)s.materials.size());
            s.materials.push_back(m);
        };
        add_sphere(0.f, -1000.f, 0.f, 1000.f, make_mat(CRT_MAT_LAMBERTIAN, 0.5f, 0.5f, 0.5f));
        for (int a = -11; a < 11; a++) {
            for (int b = -11; b < 11; b++) {
                float choose = rnd();
                float cx = (float)a + 0.9f * rnd(), cy = 0.2f, cz = (float)b + 0.9f * rnd();
                float dx = cx - 4.f, dz = cz;
                if (sqrtf(dx * dx + dz * dz) <= 0.9f) continue;
                if (choose < 0.8f) {
                    float r0 = rnd() * rnd(), g0 = rnd() * rnd(), b0 = rnd() * rnd();
                    add_sphere(cx, cy, cz, 0.2f, make_mat(CRT_MAT_LAMBERTIAN, r0, g0, b0));
                } else if (choose < 0.95f) {
                    float r0 = 0.5f + 0.5f * rnd(), g0 = 0.5f + 0.5f * rnd(), b0 = 0.5f + 0.5f * rnd();
                    float fuzz = 0.5f * rnd();
                    add_sphere(cx, cy, cz, 0.2f, make_mat(CRT_MAT_METAL, r0, g0, b0, fuzz));
                } else {
                    add_sphere(cx, cy, cz, 0.2f, make_mat(CRT_MAT_DIELECTRIC, 0.f, 0.f, 0.f, 0.f, 1.5f));
                }
            }
        }
        add_sphere(0.f, 1.f, 0.f, 1.f, make_mat(CRT_MAT_DIELECTRIC, 0.f, 0.f, 0.f, 0.f, 1.5f));
        add_sphere(-4.f, 1.f, 0.f, 1.f, make_mat(CRT_MAT_LAMBERTIAN, 0.4f, 0.2f, 0.1f));
        add_sphere(4.f, 1.f, 0.f, 1.f, make_mat(CRT_MAT_METAL, 0.7f, 0.6f, 0.5f, 0.f));
        const float from[3] = {13.f, 2.f, 3.f}, to[3] = {0.f, 0.f, 0.f};
        cam.position[0] = from[0]; cam.position[1] = from[1]; cam.position[2] = from[2];
        look_at_yaw_pitch(from, to, cam.yaw_deg, cam.pitch_deg);
        cam.vfov_deg = 20.f; cam.aspect = 1.5f; cam.aperture = 0.1f; cam.focus_dist = 10.f;
        return true;
    }
    if (name == "terrain") {
        // C4: Cornell walls + fBm height field; param = cells along x (1582 -> 5 002 284 tris).
        cornell_materials(s);
        cornell_walls(s);
        uint32_t t0 = s.num_triangles();
        int ground = append_reference_spheres(s);
        int nx = param > 0 ? param : 1582;
        int nz = nx > 1 ? nx - 1 : 1;
        add_terrain(s, nx, nz, 7u, ground);
        s.mesh_tri_offset = {0u, t0, s.num_triangles()};
        cam.position[0] = 0.f; cam.position[1] = 0.05f; cam.position[2] = 0.29f;
        cam.pitch_deg = -25.f;
        cam.vfov_deg = 80.f; cam.aspect = 16.f / 9.f; cam.focus_dist = 0.45f;
        return true;
    }
    return false;
}

}  // namespace crt
