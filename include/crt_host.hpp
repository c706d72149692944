// crt_host.hpp — C++ host side above the C-ABI (include/crt_api.h), header only.
//
// The reference's host code is C++: Raytracer owns a Camera, a SceneManager, a WindowManager and a
// CUDARenderer and calls them once per frame (reference: CudaRayTracer/src/Raytracer.h:52-106). These
// classes keep those names, method names, argument meaning and error behaviour, and forward to
// libcrt_b200.so, so that host code written against the reference reads the same here:
//
//   CRT::Camera        <- src/Core/Camera.cuh:12-211 (host half: pose, FPS controller, quality switch;
//                         the device half, getRay, lives in the library)
//   CRT::SceneManager  <- src/SceneManager.h:13-98   (scene assembly; flat arrays instead of device objects)
//   CRT::CUDARenderer  <- src/CUDARenderer.cuh:9-60  (image + RNG buffers, updateCamera, render)
//   CRT::WindowManager <- src/WindowManager.h:13-93  (headless: drawFrame = D2H copy + flipVertically into a
//                         host image, optionally written as PPM; there is no window on this box)
//   CRT::Raytracer     <- src/Raytracer.h:13-106     (run(): scripted frames instead of an event loop)
//
// Differences that are deliberate: input comes from a script (InputState per frame) instead of
// sf::Keyboard / sf::Mouse polling; errors throw std::runtime_error like CUDA_CHECK
// (src/CUDAHelpers.h:19-26) but never call cudaDeviceReset() or exit(); the function-local statics of
// Camera::updateRotation (Camera.cuh:91-96) are members, so two cameras do not share mouse state.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>

#include "crt_api.h"

namespace CRT {

static constexpr int DEFAULT_SAMPLES_PER_PIXEL = 1;          // Camera.cuh:11
static constexpr int HIGH_QUALITY_SAMPLES_PER_PIXEL = 2000;  // Camera.cuh:64

inline const char* errorName(int rc) {   // the CRT_E_* codes of crt_api.h as text, for exceptions raised before a context exists
    switch (rc) {
        case CRT_OK: return "ok";
        case CRT_E_INVALID: return "invalid argument or state";
        case CRT_E_NO_DEVICE: return "no CUDA device of compute capability 10.x (libcrt_b200 has no CPU fallback)";
        case CRT_E_CUDA: return "CUDA runtime or kernel failure";
        case CRT_E_NOMEM: return "out of memory";
        case CRT_E_IO: return "file could not be read or written";
        case CRT_E_PARSE: return "malformed OBJ / MTL";
        case CRT_E_LIMIT: return "structural limit exceeded";
        case CRT_E_BUSY: return "an asynchronous render owns the context (crt_wait)";
        default: return "unknown error";
    }
}

struct Vec3 {
    float x = 0.f, y = 0.f, z = 0.f;
    Vec3() = default;
    Vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    float length() const { return std::sqrt(x * x + y * y + z * z); }
};
inline Vec3 operator+(Vec3 a, Vec3 b) { return Vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline Vec3 operator-(Vec3 a, Vec3 b) { return Vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline Vec3 operator*(Vec3 a, float t) { return Vec3(a.x * t, a.y * t, a.z * t); }
inline Vec3 unitVector(Vec3 v) { const float inv = 1.0f / v.length(); return Vec3(inv * v.x, inv * v.y, inv * v.z); }
inline Vec3 cross(Vec3 u, Vec3 v) { return Vec3(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x); }

// What the reference polls from sf::Keyboard inside Camera::updateCamera / updatePosition
// (Camera.cuh:52,143-154), as plain data.
struct InputState {
    bool W = false, S = false, A = false, D = false, Space = false, LControl = false, F = false;
};

class Camera {
public:
    Camera() = default;
    // Camera.cuh:18-30: yaw -90, pitch 0 whatever the target is (the reference ignores `target` too)
    Camera(float aspectRatio, float fov, Vec3 position, Vec3 /*target*/, Vec3 up, float aperture, float focusDist)
        : m_Position(position), m_WorldUp(up), m_AspectRatio(aspectRatio), m_VerticalFOV(fov), m_Aperture(aperture), m_FocusDist(focusDist) {
        updateCameraVectors();
    }
    explicit Camera(const crt_camera& c)
        : m_Position(c.position[0], c.position[1], c.position[2]), m_WorldUp(c.world_up[0], c.world_up[1], c.world_up[2]),
          m_AspectRatio(c.aspect), m_VerticalFOV(c.vfov_deg), m_Aperture(c.aperture), m_FocusDist(c.focus_dist), m_Yaw(c.yaw_deg), m_Pitch(c.pitch_deg) {
        updateCameraVectors();
    }

    // Camera.cuh:46-72. Returns nothing, like the reference; read getSamplesPerPixel() afterwards.
    void updateCamera(float deltaTime, int windowWidth, int windowHeight, float mouseX, float mouseY, bool mousePressed, const InputState& keys = InputState()) {
        updateRotation(windowWidth, windowHeight, mouseX, mouseY, mousePressed);
        updatePosition(deltaTime, keys);
        updateCameraVectors();
        if (keys.F) m_HighQualityMode = !m_HighQualityMode;
        if (m_CameraRotates || m_CameraMoves) {
            m_SamplesPerPixel = DEFAULT_SAMPLES_PER_PIXEL;
            m_HighQualityMode = false;
        } else if (m_HighQualityMode) {
            m_SamplesPerPixel = HIGH_QUALITY_SAMPLES_PER_PIXEL;
        } else {
            m_SamplesPerPixel = DEFAULT_SAMPLES_PER_PIXEL;
        }
    }
    bool isCameraInMotion() const { return m_CameraRotates || m_CameraMoves; }                    // Camera.cuh:74-77
    void adjustFocusDistance(float delta) { m_FocusDist = std::fmax(0.1f, m_FocusDist + delta); }  // Camera.cuh:79-83
    float getFocusDistance() const { return m_FocusDist; }
    int getSamplesPerPixel() const { return m_SamplesPerPixel; }
    Vec3 getPosition() const { return m_Position; }
    float getYaw() const { return m_Yaw; }
    float getPitch() const { return m_Pitch; }
    void setPose(Vec3 position, float yaw, float pitch) { m_Position = position; m_Yaw = yaw; m_Pitch = pitch; updateCameraVectors(); }

    // The pose + lens the library derives its vectors from (Camera.cuh:159-182 runs inside crt_set_camera).
    crt_camera toApi() const {
        crt_camera c = {};
        c.position[0] = m_Position.x; c.position[1] = m_Position.y; c.position[2] = m_Position.z;
        c.world_up[0] = m_WorldUp.x; c.world_up[1] = m_WorldUp.y; c.world_up[2] = m_WorldUp.z;
        c.yaw_deg = m_Yaw; c.pitch_deg = m_Pitch; c.vfov_deg = m_VerticalFOV; c.aspect = m_AspectRatio;
        c.aperture = m_Aperture; c.focus_dist = m_FocusDist;
        return c;
    }

private:
    void updateRotation(int windowWidth, int windowHeight, float mouseX, float mouseY, bool mousePressed) {   // Camera.cuh:89-138
        if (!m_MouseInit) {   // the reference's function-local statics, initialised on the first call
            m_LastX = m_SmoothX = windowWidth / 2.0f;
            m_LastY = m_SmoothY = windowHeight / 2.0f;
            m_MouseInit = true;
        }
        const float smoothFactor = 0.5f;
        if (mousePressed) {
            if (m_FirstMouse) {
                m_LastX = m_SmoothX = mouseX;
                m_LastY = m_SmoothY = mouseY;
                m_FirstMouse = false;
                return;   // first frame skipped (no jump); m_CameraRotates keeps its value
            }
            m_CameraRotates = true;
            m_SmoothX = m_SmoothX * (1.0f - smoothFactor) + mouseX * smoothFactor;
            m_SmoothY = m_SmoothY * (1.0f - smoothFactor) + mouseY * smoothFactor;
            const float xoffset = (m_SmoothX - m_LastX) * -m_MouseSensitivity;
            const float yoffset = (m_SmoothY - m_LastY) * -m_MouseSensitivity;
            m_LastX = m_SmoothX;
            m_LastY = m_SmoothY;
            m_Yaw += xoffset;
            m_Pitch += yoffset;
            if (m_Pitch > 89.0f) m_Pitch = 89.0f;
            if (m_Pitch < -89.0f) m_Pitch = -89.0f;
        } else {
            m_FirstMouse = true;
            m_CameraRotates = false;
        }
    }
    void updatePosition(float deltaTime, const InputState& k) {   // Camera.cuh:139-157: moves along LAST frame's m_Front / m_Right
        const float velocity = m_MovementSpeed * deltaTime;
        const Vec3 prev = m_Position;
        if (k.W) m_Position = m_Position - m_Front * velocity;
        if (k.S) m_Position = m_Position + m_Front * velocity;
        if (k.A) m_Position = m_Position - m_Right * velocity;
        if (k.D) m_Position = m_Position + m_Right * velocity;
        if (k.Space) m_Position = m_Position + m_WorldUp * velocity;
        if (k.LControl) m_Position = m_Position - m_WorldUp * velocity;
        m_CameraMoves = m_Position.x != prev.x || m_Position.y != prev.y || m_Position.z != prev.z;
    }
    void updateCameraVectors() {   // Camera.cuh:159-169, the part the controller needs (front, right)
        const float PI = 3.1415926535897932385f;   // Core/Utility.cuh:9
        Vec3 front;
        front.x = -std::cos(m_Yaw * PI / 180.0f) * std::cos(m_Pitch * PI / 180.0f);
        front.y = -std::sin(m_Pitch * PI / 180.0f);
        front.z = -std::sin(m_Yaw * PI / 180.0f) * std::cos(m_Pitch * PI / 180.0f);
        m_Front = unitVector(front);
        m_Right = unitVector(cross(m_Front, m_WorldUp));
    }

    Vec3 m_Position, m_WorldUp = Vec3(0.f, 1.f, 0.f), m_Front, m_Right;
    float m_AspectRatio = 1.f, m_VerticalFOV = 40.f, m_Aperture = 0.f, m_FocusDist = 1.f;
    float m_Yaw = -90.0f, m_Pitch = 0.0f;
    float m_MovementSpeed = 1.0f, m_MouseSensitivity = 0.2f;
    int m_SamplesPerPixel = DEFAULT_SAMPLES_PER_PIXEL;
    bool m_HighQualityMode = false, m_CameraMoves = false, m_CameraRotates = false;
    bool m_MouseInit = false, m_FirstMouse = true;
    float m_LastX = 0.f, m_LastY = 0.f, m_SmoothX = 0.f, m_SmoothY = 0.f;
};

// Scene assembly. The reference builds device-side objects (SceneManager.h:77-98); here the scene is a
// crt_scene of flat host arrays which CUDARenderer::setScene hands to crt_set_scene.
class SceneManager {
public:
    SceneManager() = default;
    SceneManager(const SceneManager&) = delete;
    SceneManager& operator=(const SceneManager&) = delete;
    ~SceneManager() { crt_scene_free(m_Scene); }

    // one of the procedural scenes that stand in for the git-ignored assets/ ("cornell", "cornell_blob", "spheres", "terrain")
    void initializeScene(const std::string& builtin, int param = 0) {
        reset();
        if (crt_scene_builtin(builtin.c_str(), param, &m_Scene, &m_DefaultCamera) != CRT_OK)
            throw std::runtime_error("unknown built-in scene '" + builtin + "'");
        m_HasDefaultCamera = true;
    }
    // SceneManager::loadObject rules for each path (SceneManager.h:198-329), plus the two hard-coded spheres
    // of createRandomWorld (CUDAKernels.h:69-73) when addReferenceSpheres is set
    void loadObject(const std::vector<std::string>& paths, bool addReferenceSpheres = true) {
        reset();
        std::vector<const char*> p;
        for (const std::string& s : paths) p.push_back(s.c_str());
        char err[512] = {0};
        const int rc = crt_scene_load_obj(p.data(), (int32_t)p.size(), addReferenceSpheres ? 1 : 0, &m_Scene, err, sizeof(err));
        if (rc != CRT_OK) throw std::runtime_error(std::string("loadObject: ") + err);
    }
    const crt_scene* getWorld() const { return m_Scene; }
    bool hasDefaultCamera() const { return m_HasDefaultCamera; }
    const crt_camera& defaultCamera() const { return m_DefaultCamera; }

private:
    void reset() { crt_scene_free(m_Scene); m_Scene = nullptr; m_HasDefaultCamera = false; }
    crt_scene* m_Scene = nullptr;
    crt_camera m_DefaultCamera = {};
    bool m_HasDefaultCamera = false;
};

class CUDARenderer {
public:
    CUDARenderer(int width, int height, int mode = CRT_MODE_AUTO, int device = 0, bool referenceExact = false) : m_Width(width), m_Height(height) {   // CUDARenderer.cuh:31
        crt_config cfg = {};
        cfg.width = width; cfg.height = height; cfg.device = device; cfg.mode = mode;
        cfg.reference_exact = referenceExact ? 1 : 0;   // ties and unseen triangles as the reference's own BVH decides them (crt_api.h)
        const int rc = crt_create(&m_Ctx, &cfg);
        if (rc != CRT_OK) {
            // crt_create leaves *out_ctx null on failure; its detail message is kept under the null context
            std::string why = crt_last_error(m_Ctx);
            if (why.empty()) why = errorName(rc);
            crt_destroy(m_Ctx);
            m_Ctx = nullptr;
            throw std::runtime_error("CUDA error: " + why + " (code " + std::to_string(rc) + ")");
        }
    }
    CUDARenderer(const CUDARenderer&) = delete;
    CUDARenderer& operator=(const CUDARenderer&) = delete;
    ~CUDARenderer() { crt_destroy(m_Ctx); }                                                       // :34-37

    // :39-49; the reference seeds with rand() (first glibc value 1804289383)
    void initialize(unsigned long long seed = 1804289383ull, unsigned long long subsequenceBase = 0) { check(crt_init_rng(m_Ctx, seed, subsequenceBase)); }
    void setScene(const SceneManager& scene) {   // SceneManager::initializeScene's device half
        if (!scene.getWorld()) throw std::runtime_error("setScene: the SceneManager holds no scene");
        check(crt_set_scene(m_Ctx, scene.getWorld()));
    }
    void updateCamera(const Camera& camera) { const crt_camera c = camera.toApi(); check(crt_set_camera(m_Ctx, &c)); }   // :51-53
    // :55-60. Samples per pixel travel with the call instead of inside the camera constant (Camera.cuh:185);
    // MAX_BOUNCES / MIN_BOUNCES / MAX_PROB (CUDAKernels.h:106-108) are parameters with the reference's values.
    void render(int samplesPerPixel, int maxBounces = 20, unsigned flags = 0) {
        crt_render_params p = {};
        p.spp = samplesPerPixel; p.max_bounces = maxBounces; p.min_bounces = 3; p.rr_max_prob = 0.95f; p.flags = flags;
        check(crt_render(m_Ctx, &p));
    }
    unsigned char* getImageData() const {   // :17 — DEVICE pointer, RGBA8, row 0 = bottom
        void* dev = nullptr;
        check(crt_get_image_rgba8(m_Ctx, nullptr, &dev));
        return static_cast<unsigned char*>(dev);
    }
    void copyImageToHost(unsigned char* hostRgba8) const { check(crt_get_image_rgba8(m_Ctx, hostRgba8, nullptr)); }
    void copyAccumToHost(float* hostRgba32f) const { check(crt_get_accum(m_Ctx, hostRgba32f, nullptr)); }
    void* getRandState() const { void* dev = nullptr; check(crt_get_rng_state(m_Ctx, nullptr, &dev)); return dev; }   // :18 (6 planes of u32)
    crt_stats stats() const { crt_stats s = {}; check(crt_get_stats(m_Ctx, &s)); return s; }
    void writePPM(const std::string& path) const { check(crt_write_ppm(m_Ctx, path.c_str())); }
    void writePFM(const std::string& path) const { check(crt_write_pfm(m_Ctx, path.c_str())); }
    int width() const { return m_Width; }
    int height() const { return m_Height; }
    crt_ctx* context() const { return m_Ctx; }

private:
    void check(int rc) const {   // CUDA_CHECK's convention (CUDAHelpers.h:19-26) without the device reset
        if (rc != CRT_OK) throw std::runtime_error(std::string("CUDA error: ") + crt_last_error(m_Ctx) + " (code " + std::to_string(rc) + ")");
    }
    crt_ctx* m_Ctx = nullptr;
    int m_Width, m_Height;
};

// CUDARenderer over several GPUs of this process (crt_group_*): same calls; a frame's samples are sharded
// over the devices and the root device (devices[0]) ends up with the summed buffer and the image. The reference has
// nothing like it (one device: CUDARenderer.cuh:55-60).
class CUDARendererGroup {
public:
    CUDARendererGroup(int width, int height, const std::vector<int>& devices, int mode = CRT_MODE_AUTO) : m_Width(width), m_Height(height) {
        crt_config cfg = {};
        cfg.width = width; cfg.height = height; cfg.mode = mode;
        std::vector<int32_t> d(devices.begin(), devices.end());
        const int rc = crt_group_create(&m_Group, &cfg, d.data(), (int32_t)d.size());
        if (rc != CRT_OK) {
            std::string why = crt_last_error(nullptr);
            if (why.empty()) why = errorName(rc);
            throw std::runtime_error("CUDA error: " + why + " (code " + std::to_string(rc) + ")");
        }
    }
    CUDARendererGroup(const CUDARendererGroup&) = delete;
    CUDARendererGroup& operator=(const CUDARendererGroup&) = delete;
    ~CUDARendererGroup() { crt_group_destroy(m_Group); }
    void initialize(unsigned long long seed = 1804289383ull) { check(crt_group_init_rng(m_Group, seed)); }
    void setScene(const SceneManager& scene) {
        if (!scene.getWorld()) throw std::runtime_error("setScene: the SceneManager holds no scene");
        check(crt_group_set_scene(m_Group, scene.getWorld()));
    }
    void updateCamera(const Camera& camera) { const crt_camera c = camera.toApi(); check(crt_group_set_camera(m_Group, &c)); }
    void render(int samplesPerPixel, int maxBounces = 20, unsigned flags = 0, unsigned groupFlags = 0) {
        crt_render_params p = {};
        p.spp = samplesPerPixel; p.max_bounces = maxBounces; p.min_bounces = 3; p.rr_max_prob = 0.95f; p.flags = flags;
        check(crt_group_render(m_Group, &p, groupFlags));
    }
    void copyImageToHost(unsigned char* hostRgba8) const { checkRoot(crt_get_image_rgba8(root(), hostRgba8, nullptr)); }
    void copyAccumToHost(float* hostRgba32f) const { checkRoot(crt_get_accum(root(), hostRgba32f, nullptr)); }
    crt_stats stats(float* reduceMs = nullptr, bool* peer = nullptr) const {
        crt_stats s = {};
        int32_t p = 0;
        check(crt_group_get_stats(m_Group, &s, reduceMs, &p));
        if (peer) *peer = p != 0;
        return s;
    }
    crt_stats rootStats() const { crt_stats s = {}; checkRoot(crt_get_stats(root(), &s)); return s; }
    void writePPM(const std::string& path) const { checkRoot(crt_write_ppm(root(), path.c_str())); }
    void writePFM(const std::string& path) const { checkRoot(crt_write_pfm(root(), path.c_str())); }
    int size() const { return crt_group_size(m_Group); }
    crt_ctx* root() const { return crt_group_ctx(m_Group, 0); }

private:
    void check(int rc) const {
        if (rc != CRT_OK) throw std::runtime_error(std::string("CUDA error: ") + crt_group_last_error(m_Group) + " (code " + std::to_string(rc) + ")");
    }
    void checkRoot(int rc) const {
        if (rc != CRT_OK) throw std::runtime_error(std::string("CUDA error: ") + crt_last_error(root()) + " (code " + std::to_string(rc) + ")");
    }
    crt_group* m_Group = nullptr;
    int m_Width, m_Height;
};

// WindowManager::drawFrame without the window (WindowManager.h:79-93): device image -> host, flipVertically.
class WindowManager {
public:
    WindowManager(int width, int height) : m_Width(width), m_Height(height), m_Image((size_t)width * height * 4) {}
    void drawFrame(const CUDARenderer& renderer) {
        std::vector<unsigned char> bottomUp(m_Image.size());
        renderer.copyImageToHost(bottomUp.data());
        const size_t row = (size_t)m_Width * 4;
        for (int y = 0; y < m_Height; y++)   // sf::Image::flipVertically
            std::copy(bottomUp.begin() + (size_t)(m_Height - 1 - y) * row, bottomUp.begin() + (size_t)(m_Height - y) * row, m_Image.begin() + (size_t)y * row);
        m_Frames++;
    }
    const std::vector<unsigned char>& image() const { return m_Image; }   // top-down RGBA8, what the window would show
    int frames() const { return m_Frames; }

private:
    int m_Width, m_Height, m_Frames = 0;
    std::vector<unsigned char> m_Image;
};

// One scripted frame of Raytracer::run: what handleEvents + the mouse / keyboard polls would have produced.
struct FrameInput {
    float deltaTime = 1.0f / 60.0f;
    InputState keys;
    float mouseX = 0.f, mouseY = 0.f;
    bool rightMousePressed = false;
    float focusDelta = 0.f;   // PageUp / PageDown: +-0.1 (WindowManager.h:62-71)
};

struct FrameRecord {
    int samplesPerPixel;
    float renderMs;
    unsigned long long rays;
};

class Raytracer {
public:
    // Raytracer.h:36-46 + initializeScene :77-92 (camera at (0,4,4) in the reference; here the scene's own
    // default camera when it has one)
    Raytracer(int width, int height, float verticalFOV, float aperture, const std::string& builtinScene, int sceneParam = 0, int mode = CRT_MODE_AUTO)
        : m_Width(width), m_Height(height), m_WindowManager(width, height), m_Renderer(width, height, mode) {
        m_SceneManager.initializeScene(builtinScene, sceneParam);
        if (m_SceneManager.hasDefaultCamera()) {
            crt_camera c = m_SceneManager.defaultCamera();
            c.aspect = (float)width / (float)height;
            m_Camera = Camera(c);
        } else {
            const Vec3 position(0, 4, 4), target(0, 0, 0);
            m_Camera = Camera((float)width / (float)height, verticalFOV, position, target, Vec3(0, 1, 0), aperture, (position - target).length());
        }
        m_Renderer.initialize();
        m_Renderer.setScene(m_SceneManager);
    }

    // Raytracer::run (:52-71) over scripted frames; highQualitySpp caps the reference's 2000-spp still frames
    std::vector<FrameRecord> run(const std::vector<FrameInput>& script, int maxBounces = 20, int highQualitySpp = HIGH_QUALITY_SAMPLES_PER_PIXEL, bool draw = true) {
        std::vector<FrameRecord> out;
        for (const FrameInput& in : script) {
            if (in.focusDelta != 0.f) m_Camera.adjustFocusDistance(in.focusDelta);   // handleEvents
            updateAndRender(in, maxBounces, highQualitySpp);
            if (draw) drawFrame();
            const crt_stats st = m_Renderer.stats();
            out.push_back(FrameRecord{m_LastSpp, st.render_ms, (unsigned long long)st.rays});
        }
        return out;
    }
    Camera& camera() { return m_Camera; }
    CUDARenderer& renderer() { return m_Renderer; }
    WindowManager& window() { return m_WindowManager; }

private:
    void updateAndRender(const FrameInput& in, int maxBounces, int highQualitySpp) {   // :94-102
        m_Camera.updateCamera(in.deltaTime, m_Width, m_Height, in.mouseX, in.mouseY, in.rightMousePressed, in.keys);
        m_Renderer.updateCamera(m_Camera);
        m_LastSpp = m_Camera.getSamplesPerPixel();
        if (m_LastSpp == HIGH_QUALITY_SAMPLES_PER_PIXEL) m_LastSpp = highQualitySpp;
        m_Renderer.render(m_LastSpp, maxBounces);
    }
    void drawFrame() { m_WindowManager.drawFrame(m_Renderer); }                        // :104-106

    int m_Width, m_Height, m_LastSpp = DEFAULT_SAMPLES_PER_PIXEL;
    Camera m_Camera;
    SceneManager m_SceneManager;
    WindowManager m_WindowManager;
    CUDARenderer m_Renderer;
};

}  // namespace CRT
