// crt_cli.cpp — headless stand-in for the reference's application (reference: CudaRayTracer/src/EntryPoint.cu:7-44:
// constants, Raytracer raytracer(...); raytracer.run(); errors caught in main). Links against libcrt_b200.so through
// include/crt_host.hpp; no rendering code lives here.
//
//   crt_cli --scene cornell_blob --width 2560 --height 1440 --spp 2000 --bounces 20 --ppm out.ppm
//   crt_cli --scene cornell_blob --width 1920 --height 1080 --frames 200 --orbit 0.5      progressive, 1 spp per frame while moving
//   crt_cli --controller-trace script.txt --width 800 --height 600                        camera controller only (no GPU needed)
//
// script lines: dt keys mouse_x mouse_y right_pressed focus_delta      (keys: letters of W S A D U(p=Space) C(trl) F, or '-')
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>

#include "../../include/crt_host.hpp"

namespace {

struct Options {
    std::string scene = "cornell", obj, ppm, pfm, trace;
    int param = 0, width = 800, height = 600, spp = 0, bounces = 20, frames = 1, gpus = 1, force_nccl = 0, ref_exact = 0, mode = CRT_MODE_AUTO, hq_spp = CRT::HIGH_QUALITY_SAMPLES_PER_PIXEL;
    unsigned long long seed = 1804289383ull;
    float orbit = 0.f, vfov = 90.f, aperture = 0.f;   // EntryPoint.cu:11-12 defaults
};

void usage() {
    std::printf("usage: crt_cli [--scene cornell|cornell_blob|spheres|terrain] [--param N] [--obj a.obj[,b.obj]]\n"
                "               [--width W] [--height H] [--spp N] [--bounces B] [--frames F] [--orbit DEG_PER_FRAME]\n"
                "               [--mode wavefront|megakernel|auto] [--seed S] [--hq-spp N] [--ppm FILE] [--pfm FILE]\n"
                "               [--controller-trace SCRIPT] [--gpus N] [--nccl] [--reference-exact]\n"
                "  --spp N            one still frame of N samples per pixel (the reference's high-quality frame)\n"
                "  --frames F         F progressive frames through Raytracer::run (1 spp while the camera moves)\n"
                "  --gpus N           with --spp: shard the frame's samples over GPUs 0..N-1 of this process (crt_group_*)\n"
                "  --nccl             with --gpus: reduce with ncclReduce instead of the peer-memory kernel\n"
                "  --reference-exact  equal-t ties and flat-leaf triangles as the reference's own BVH treats them (slower scene set-up)\n"
                "  --controller-trace print the camera pose after every scripted frame; host logic only, no GPU\n");
}

CRT::InputState parse_keys(const std::string& k) {
    CRT::InputState s;
    for (char c : k) {
        if (c == 'W') s.W = true; else if (c == 'S') s.S = true; else if (c == 'A') s.A = true; else if (c == 'D') s.D = true;
        else if (c == 'U') s.Space = true; else if (c == 'C') s.LControl = true; else if (c == 'F') s.F = true;
    }
    return s;
}

std::vector<CRT::FrameInput> read_script(const std::string& path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open script " + path);
    std::vector<CRT::FrameInput> out;
    std::string line;
    while (std::getline(f, line)) {
        if (line.empty() || line[0] == '#') continue;
        std::istringstream is(line);
        CRT::FrameInput in;
        std::string keys;
        int pressed = 0;
        if (!(is >> in.deltaTime >> keys >> in.mouseX >> in.mouseY >> pressed >> in.focusDelta)) throw std::runtime_error("bad script line: " + line);
        in.keys = parse_keys(keys);
        in.rightMousePressed = pressed != 0;
        out.push_back(in);
    }
    return out;
}

int controller_trace(const Options& o) {
    // same starting pose as the reference application: (0,4,4) looking down -z (Raytracer.h:79-84)
    const CRT::Vec3 position(0, 4, 4), target(0, 0, 0);
    CRT::Camera cam((float)o.width / (float)o.height, o.vfov, position, target, CRT::Vec3(0, 1, 0), o.aperture, (position - target).length());
    for (const CRT::FrameInput& in : read_script(o.trace)) {
        if (in.focusDelta != 0.f) cam.adjustFocusDistance(in.focusDelta);
        cam.updateCamera(in.deltaTime, o.width, o.height, in.mouseX, in.mouseY, in.rightMousePressed, in.keys);
        const CRT::Vec3 p = cam.getPosition();
        std::printf("%.9g %.9g %.9g %.9g %.9g %.9g %d %d\n", p.x, p.y, p.z, cam.getYaw(), cam.getPitch(), cam.getFocusDistance(),
                    cam.getSamplesPerPixel(), cam.isCameraInMotion() ? 1 : 0);
    }
    return 0;
}

int run(const Options& o) {
    if (!o.trace.empty()) return controller_trace(o);
    CRT::SceneManager scene;
    if (!o.obj.empty()) {
        std::vector<std::string> paths;
        std::stringstream ss(o.obj);
        for (std::string p; std::getline(ss, p, ',');) paths.push_back(p);
        scene.loadObject(paths);
    } else {
        scene.initializeScene(o.scene, o.param);
    }
    CRT::Camera camera;
    if (scene.hasDefaultCamera()) {
        crt_camera c = scene.defaultCamera();
        c.aspect = (float)o.width / (float)o.height;
        camera = CRT::Camera(c);
    } else {
        const CRT::Vec3 position(0, 4, 4), target(0, 0, 0);
        camera = CRT::Camera((float)o.width / (float)o.height, o.vfov, position, target, CRT::Vec3(0, 1, 0), o.aperture, (position - target).length());
    }
    if (o.gpus > 1) {   // one still frame over several GPUs: the spp budget sharded, one reduce onto GPU 0
        if (o.spp <= 0) throw std::runtime_error("--gpus needs --spp (a still frame); progressive 1-spp frames do not shard");
        std::vector<int> devices;
        for (int d = 0; d < o.gpus; d++) devices.push_back(d);
        CRT::CUDARendererGroup group(o.width, o.height, devices, o.mode);
        group.initialize(o.seed);
        group.setScene(scene);
        group.updateCamera(camera);
        group.render(o.spp, o.bounces, 0, o.force_nccl ? CRT_GROUP_FORCE_NCCL : 0);
        float reduce_ms = 0.f;
        bool peer = false;
        const crt_stats st = group.stats(&reduce_ms, &peer);
        if (!o.ppm.empty()) group.writePPM(o.ppm);
        if (!o.pfm.empty()) group.writePFM(o.pfm);
        std::printf("{\"frames\": 1, \"gpus\": %d, \"width\": %d, \"height\": %d, \"rays\": %llu, \"render_ms\": %.3f, \"reduce_ms\": %.3f, "
                    "\"reduce\": \"%s\", \"mrays_per_s\": %.1f}\n", o.gpus, o.width, o.height, (unsigned long long)st.rays, st.render_ms, reduce_ms,
                    peer && !o.force_nccl ? "peer-memory kernel" : "ncclReduce", st.render_ms > 0 ? st.rays / (st.render_ms + reduce_ms) / 1e3 : 0.0);
        return 0;
    }
    CRT::CUDARenderer renderer(o.width, o.height, o.mode, 0, o.ref_exact != 0);
    CRT::WindowManager window(o.width, o.height);
    renderer.initialize(o.seed);
    renderer.setScene(scene);
    const crt_stats build = renderer.stats();
    double total_ms = 0.0;
    unsigned long long total_rays = 0;
    int frames = 0;
    if (o.spp > 0) {   // one still frame
        renderer.updateCamera(camera);
        renderer.render(o.spp, o.bounces);
        window.drawFrame(renderer);
        const crt_stats st = renderer.stats();
        total_ms = st.render_ms; total_rays = st.rays; frames = 1;
    } else {           // progressive frames; --orbit turns the camera by dragging the mouse a fixed distance per frame
        float mouse_x = o.width / 2.0f;
        for (int f = 0; f < o.frames; f++) {
            const bool drag = o.orbit != 0.f;
            if (drag) mouse_x += o.orbit / 0.2f * 2.0f;   // sensitivity 0.2 deg per pixel, smoothing 0.5
            camera.updateCamera(1.0f / 60.0f, o.width, o.height, mouse_x, o.height / 2.0f, drag);
            renderer.updateCamera(camera);
            int spp = camera.getSamplesPerPixel();
            if (spp == CRT::HIGH_QUALITY_SAMPLES_PER_PIXEL) spp = o.hq_spp;
            renderer.render(spp, o.bounces);
            window.drawFrame(renderer);
            const crt_stats st = renderer.stats();
            total_ms += st.render_ms; total_rays += st.rays; frames++;
        }
    }
    if (!o.ppm.empty()) renderer.writePPM(o.ppm);
    if (!o.pfm.empty()) renderer.writePFM(o.pfm);
    std::printf("{\"frames\": %d, \"width\": %d, \"height\": %d, \"rays\": %llu, \"render_ms\": %.3f, \"mrays_per_s\": %.1f, "
                "\"bvh_nodes\": %u, \"bvh_build_ms\": %.3f, \"rng_init_ms\": %.3f}\n",
                frames, o.width, o.height, total_rays, total_ms, total_ms > 0 ? total_rays / total_ms / 1e3 : 0.0, build.bvh_nodes, build.bvh_build_ms,
                build.rng_init_ms);
    return 0;
}

}  // namespace

int main(int argc, char** argv) {
    Options o;
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&]() -> const char* { if (i + 1 >= argc) { std::fprintf(stderr, "missing value after %s\n", a.c_str()); std::exit(2); } return argv[++i]; };
        if (a == "--help" || a == "-h") { usage(); return 0; }
        else if (a == "--scene") o.scene = next();
        else if (a == "--param") o.param = std::atoi(next());
        else if (a == "--obj") o.obj = next();
        else if (a == "--width") o.width = std::atoi(next());
        else if (a == "--height") o.height = std::atoi(next());
        else if (a == "--spp") o.spp = std::atoi(next());
        else if (a == "--bounces") o.bounces = std::atoi(next());
        else if (a == "--frames") o.frames = std::atoi(next());
        else if (a == "--orbit") o.orbit = (float)std::atof(next());
        else if (a == "--seed") o.seed = std::strtoull(next(), nullptr, 10);
        else if (a == "--hq-spp") o.hq_spp = std::atoi(next());
        else if (a == "--ppm") o.ppm = next();
        else if (a == "--pfm") o.pfm = next();
        else if (a == "--controller-trace") o.trace = next();
        else if (a == "--gpus") o.gpus = std::atoi(next());
        else if (a == "--nccl") o.force_nccl = 1;
        else if (a == "--reference-exact") o.ref_exact = 1;
        else if (a == "--mode") { const std::string m = next(); o.mode = m == "wavefront" ? CRT_MODE_WAVEFRONT : m == "megakernel" ? CRT_MODE_MEGAKERNEL : CRT_MODE_AUTO; }
        else { std::fprintf(stderr, "unknown option %s\n", a.c_str()); usage(); return 2; }
    }
    try {
        return run(o);
    } catch (const std::exception& e) {   // EntryPoint.cu:38-41
        std::fprintf(stderr, "Error: %s\n", e.what());
        return 1;
    }
}
