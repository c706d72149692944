/*
 * crt_api.h — C-ABI of the B200-native render path (libcrt_b200.so).
 *
 * This is the drop-in boundary for the render path of Mordentary/RayTracer-Cuda.
 * The reference has no FFI layer; its de-facto operator interface is the C++ class
 * `CUDARenderer` plus the scene it is fed by `SceneManager`
 * (reference: CudaRayTracer/src/CUDARenderer.cuh:9-60, src/SceneManager.h:19-98).
 * Every entry point below names the reference interface it replaces.
 *
 * Conventions
 *  - plain C: pointers + sizes, no C++/torch types; never throws, never exits,
 *    never calls cudaDeviceReset (reference: src/CUDAHelpers.h:19-26 does all three).
 *  - every function returns CRT_OK (0) or a negative CRT_E_* code;
 *    crt_last_error(ctx) returns the message of the last failure on that context.
 *  - host input arrays are copied; the caller keeps ownership.
 *  - device pointers handed out are borrowed, valid until the next crt_set_scene /
 *    crt_destroy on that context.
 *  - one context = one device = one host thread at a time. No global symbols
 *    (the reference's `__constant__ Camera d_camera`, src/Core/Camera.cuh:213, made it
 *    one-renderer-per-process); several contexts may coexist.
 *  - there is NO CPU fallback: without a CUDA device every call that needs one fails
 *    with CRT_E_NO_DEVICE.
 *  - image rows: row 0 is the BOTTOM row, exactly like the reference's device image
 *    (src/CUDAKernels.h:152, flipped on display by src/WindowManager.h:89).
 */
#ifndef CRT_API_H
#define CRT_API_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRT_API_VERSION 2

/* ---- error codes -------------------------------------------------------- */
enum {
    CRT_OK            =  0,
    CRT_E_INVALID     = -1,  /* bad argument / bad state (e.g. render before set_scene) */
    CRT_E_NO_DEVICE   = -2,  /* no CUDA device / wrong architecture (needs sm_100)       */
    CRT_E_CUDA        = -3,  /* a CUDA runtime call or kernel failed; see crt_last_error */
    CRT_E_NOMEM       = -4,
    CRT_E_IO          = -5,  /* file could not be read / written                         */
    CRT_E_PARSE       = -6,  /* malformed OBJ / MTL                                      */
    CRT_E_LIMIT       = -7,  /* a structural limit was exceeded (stack depth, counts)    */
    CRT_E_BUSY        = -8   /* an asynchronous render owns the context: call crt_wait   */
};

/* ---- materials: mirrors CRT::MaterialType / CRT::MaterialData ------------
 * reference: src/Core/Material.cuh:8-47 (same enum order, same 36-byte POD). */
enum {
    CRT_MAT_LAMBERTIAN    = 0,
    CRT_MAT_METAL         = 1,
    CRT_MAT_DIELECTRIC    = 2,
    CRT_MAT_DIFFUSE_LIGHT = 3
};

typedef struct crt_material {
    int32_t type;          /* CRT_MAT_*                                            */
    float   albedo[3];     /* Lambertian / Metal attenuation; finite and >= 0, else
                              crt_set_scene fails with CRT_E_INVALID (-0.0 is stored as +0.0) */
    float   emission[3];   /* DiffuseLight radiance                                 */
    float   roughness;     /* Metal fuzz; clamped from above to 1 (Material.cuh:86) */
    float   ior;           /* Dielectric index of refraction                        */
} crt_material;

/* ---- scene: flat host arrays ---------------------------------------------
 * replaces SceneManager::initializeScene + the device-side object construction
 * kernels createRandomWorld / initMesh / createBVH
 * (reference: src/SceneManager.h:77-196, src/CUDAKernels.h:28-100).
 *
 * Primitive id (what crt_primary_ids reports): triangle i of `indices` has id i;
 * sphere j has id num_triangles + j; a miss is -1.
 * `mesh_tri_offset` (num_meshes+1 entries, may be NULL = one mesh) only records how the
 * reference would group triangles into CRT::Mesh objects; this library ignores it for
 * rendering (one flat BVH) but the oracle and the reference shim use it. */
typedef struct crt_scene {
    const float*    positions;        /* xyz per vertex, num_vertices*3 floats          */
    uint32_t        num_vertices;
    const uint32_t* indices;          /* 3 per triangle (global vertex numbers)         */
    const int32_t*  tri_material;     /* material index per triangle                    */
    uint32_t        num_triangles;
    const float*    spheres;          /* cx,cy,cz,r per sphere                          */
    const int32_t*  sphere_material;
    uint32_t        num_spheres;
    const crt_material* materials;
    uint32_t        num_materials;
    const uint32_t* mesh_tri_offset;  /* optional, num_meshes+1 entries                 */
    uint32_t        num_meshes;
} crt_scene;

/* ---- camera: mirrors CRT::Camera's constructor + yaw/pitch state ---------
 * reference: src/Core/Camera.cuh:18-30 (ctor), :159-182 (updateCameraVectors),
 * :188-202 (fields). View direction is -front(yaw,pitch); yaw=-90,pitch=0 looks
 * along -z. Angles in degrees. */
typedef struct crt_camera {
    float position[3];
    float world_up[3];
    float yaw_deg;
    float pitch_deg;
    float vfov_deg;
    float aspect;
    float aperture;
    float focus_dist;
} crt_camera;

/* ---- context configuration ------------------------------------------------
 * replaces CUDARenderer(int w,int h) + createRenderConfig
 * (reference: src/CUDARenderer.cuh:31-32, src/CUDAHelpers.h:28-35). */
enum {
    CRT_MODE_WAVEFRONT  = 0,  /* raygen / extend / shade / accumulate kernels (default) */
    CRT_MODE_MEGAKERNEL = 1,  /* one thread per pixel, spp loop inside, for comparison  */
    CRT_MODE_AUTO       = 2   /* wavefront for large frames over non-trivial scenes (>= 1.5 M pixels and
                                 >= 4096 primitives), megakernel otherwise; same bits either way */
};

typedef struct crt_config {
    int32_t width;
    int32_t height;
    int32_t device;        /* CUDA device ordinal                                       */
    int32_t mode;          /* CRT_MODE_*                                                */
    int32_t bvh_leaf_max;  /* 0 = default (2), 1..3                                        */
    int32_t reference_exact; /* 1 = follow the reference where its results depend on ITS acceleration structure
                              (see crt_scene_reference_order): equal-t ties go to the primitive the reference
                              tests last, triangles in zero-thickness leaves of its mesh BVH are not seen.
                              crt_set_scene then replays the reference's sequential partition on the host
                              (~5 s per million triangles). 0 (default): larger primitive id wins a tie,
                              every triangle is seen                                     */
    int32_t reserved[6];   /* must be zero                                              */
} crt_config;

/* ---- per-render parameters -------------------------------------------------
 * The reference hard-codes MAX_BOUNCES=20, MIN_BOUNCES=3, MAX_PROB=0.95
 * (src/CUDAKernels.h:106-108) and keeps spp in the camera (src/Core/Camera.cuh:185). */
enum {
    CRT_RENDER_ACCUMULATE = 1u << 0, /* add to the float4 accumulation buffer instead of
                                        overwriting it (the reference always overwrites,
                                        src/CUDAKernels.h:164)                          */
    CRT_RENDER_NO_TONEMAP = 1u << 1, /* skip the RGBA8 pack (multi-GPU ranks != 0)      */
    CRT_RENDER_ASYNC = 1u << 3,      /* crt_render returns at once; the frame is driven by a host thread of the
                                        library. Until crt_wait(ctx) returns the frame's status, every other call
                                        on the context fails with CRT_E_BUSY (crt_get_stats and crt_last_error
                                        excepted)                                                              */
    CRT_RENDER_TIME_KERNELS = 1u << 2 /* bracket every kernel launch with CUDA events on the
                                        launching stream and report per-kernel-class totals in
                                        crt_stats (extend_ms / shade_ms); off by default       */
};

typedef struct crt_render_params {
    int32_t  spp;           /* samples per pixel rendered by this call                   */
    int32_t  max_bounces;   /* reference default 20                                      */
    int32_t  min_bounces;   /* Russian roulette starts at this bounce; reference 3       */
    float    rr_max_prob;   /* reference 0.95                                            */
    uint32_t flags;         /* CRT_RENDER_*                                              */
    int32_t  reserved[3];   /* must be zero                                              */
} crt_render_params;

/* ---- statistics ------------------------------------------------------------ */
typedef struct crt_stats {
    uint64_t rays;             /* closest-hit queries traced by the last crt_render       */
    uint64_t rays_total;       /* since crt_create                                        */
    uint64_t samples;          /* camera samples started by the last crt_render           */
    float    render_ms;        /* device time of the last crt_render (CUDA events)        */
    float    bvh_build_ms;     /* device time of the last BVH build                       */
    float    rng_init_ms;      /* device time of the last crt_init_rng                    */
    float    scene_upload_ms;  /* host->device copies in crt_set_scene                    */
    uint32_t bvh_nodes;        /* wide-BVH node count                                     */
    uint32_t bvh_prims;        /* primitive records in leaf order                         */
    uint32_t bvh_depth;        /* wide-BVH depth                                          */
    uint32_t kernel_launches;  /* kernels launched by the last crt_render                 */
    /* traversal work counters, filled only by crt_render_counted */
    uint64_t nodes_fetched;
    uint64_t prims_tested;
    /* per-kernel-class device time of the last crt_render, filled with CRT_RENDER_TIME_KERNELS:
     * wavefront: extend_ms = sum over k_wf_extend launches, shade_ms = sum over k_wf_shade launches;
     * megakernel: extend_ms = the one k_render_mega launch, shade_ms = 0 */
    float    extend_ms;
    float    shade_ms;
    uint32_t extend_launches;
    uint32_t shade_launches;
    uint32_t bvh_giants;       /* primitives kept out of the BVH and tested by every ray (<= 4)  */
    float    tail_ms;          /* wavefront, with CRT_RENDER_TIME_KERNELS: the closing extend + k_wf_tail launches   */
    float    begin_ms;         /* wavefront, with CRT_RENDER_TIME_KERNELS: k_wf_begin                                 */
    float    compact_ms;       /* wavefront, with CRT_RENDER_TIME_KERNELS: the k_compact_* triples                    */
    uint32_t host_syncs;       /* blocking stream synchronisations inside the last crt_render (queue-length read-backs) */
    uint32_t iterations;       /* wavefront iterations (extend + shade per part) of the last crt_render               */
    uint32_t graph_launches;   /* batches of the last crt_render replayed as one CUDA graph launch (their kernels are
                                  counted in kernel_launches all the same)                                             */
    uint32_t unseen_triangles; /* reference-exact mode: triangles the reference's zero-thickness leaves hide (made degenerate here) */
} crt_stats;

typedef struct crt_ctx crt_ctx;

/* ---- lifecycle --------------------------------------------------------------- */

/* replaces CUDARenderer::CUDARenderer + the cudaMalloc half of CUDARenderer::initialize
 * (reference: src/CUDARenderer.cuh:31-32, :39-43). */
int crt_create(crt_ctx** out_ctx, const crt_config* cfg);

/* replaces CUDARenderer::~CUDARenderer + SceneManager::~SceneManager
 * (reference: src/CUDARenderer.cuh:34-37, src/SceneManager.h:66-75). NULL is a no-op. */
void crt_destroy(crt_ctx* ctx);

/* Message of the last failure on ctx (or of crt_create when ctx is NULL). Never NULL. */
const char* crt_last_error(const crt_ctx* ctx);

int crt_api_version(void);

/* ---- RNG ------------------------------------------------------------------------
 * replaces the initRandState launch of CUDARenderer::initialize
 * (reference: src/CUDARenderer.cuh:46, src/CUDAKernels.h:18-26):
 *   state[pixel] = curand_init(seed, subsequence_base + pixel, 0)   (XORWOW)
 * The reference passes seed = rand() (1804289383 with glibc) and base 0.
 * Multi-GPU rank g uses subsequence_base = g*width*height (disjoint streams). */
int crt_init_rng(crt_ctx* ctx, uint64_t seed, uint64_t subsequence_base);

/* ---- scene + camera ---------------------------------------------------------- */

/* replaces SceneManager::initializeScene (reference: src/SceneManager.h:77-98):
 * uploads the flat arrays, builds the wide BVH on the GPU. */
int crt_set_scene(crt_ctx* ctx, const crt_scene* scene);

/* replaces CUDARenderer::updateCamera (reference: src/CUDARenderer.cuh:51-53) and the
 * host-side Camera::updateCameraVectors (src/Core/Camera.cuh:159-182). */
int crt_set_camera(crt_ctx* ctx, const crt_camera* cam);

/* ---- render --------------------------------------------------------------------
 * replaces CUDARenderer::render (reference: src/CUDARenderer.cuh:55-60 →
 * CUDAKernels::render, src/CUDAKernels.h:147-166). Synchronous on return.
 * Continues each pixel's RNG stream where the previous call left it, like the
 * reference (src/CUDAKernels.h:153,165). */
int crt_render(crt_ctx* ctx, const crt_render_params* params);

/* Joins an asynchronous render (CRT_RENDER_ASYNC) and returns its status; CRT_OK at once when none is in flight. */
int crt_wait(crt_ctx* ctx);

/* Same as crt_render but with per-ray node/primitive counters switched on (slower);
 * fills crt_stats.nodes_fetched / prims_tested. Used for the roofline bytes-per-ray. */
int crt_render_counted(crt_ctx* ctx, const crt_render_params* params);

/* Primary-hit primitive-id buffer for parity: traces sample 0's camera ray of every
 * pixel from a FRESH stream curand_init(seed, subsequence_base+pixel, 0) (it does not
 * disturb the context's RNG state). out_ids: width*height int32 (host), out_t: optional
 * width*height float hit parameter (host, may be NULL; +inf on a miss). */
int crt_primary_ids(crt_ctx* ctx, uint64_t seed, uint64_t subsequence_base,
                    int32_t* out_ids, float* out_t);

/* The same buffers as crt_primary_ids (same kernel, bit-identical output) WITHOUT any allocation: the id / t staging
 * that crt_primary_ids takes from cudaMalloc and frees again lives in a file-scope device array of the library
 * (CRT_PRIMARY_STATIC_MAX_PIXELS * 8 bytes = 8 MB, held by every GPU that loads the library until the process
 * exits). For callers that must not reach cudaMalloc / cudaFree: tight serving loops, harnesses that forbid it.
 * This entry point has NO counterpart in the reference.
 * width*height above CRT_PRIMARY_STATIC_MAX_PIXELS fails with CRT_E_LIMIT (use crt_primary_ids). All calls on a GPU
 * share the one array: it keeps the last call's contents (every pixel of the frame is rewritten, nothing needs
 * clearing), and calls that overlap on the same GPU (two contexts, two host threads) race on it. */
#define CRT_PRIMARY_STATIC_MAX_PIXELS (1u << 20)
int crt_primary_ids_static(crt_ctx* ctx, uint64_t seed, uint64_t subsequence_base,
                           int32_t* out_ids, float* out_t);

/* ---- results -------------------------------------------------------------------- */

/* replaces CUDARenderer::getImageData (reference: src/CUDARenderer.cuh:17) + the D2H
 * copy of WindowManager::drawFrame (src/WindowManager.h:88). host_rgba8 may be NULL;
 * out_dev (may be NULL) receives the borrowed device pointer. width*height*4 bytes,
 * row 0 = bottom. */
int crt_get_image_rgba8(crt_ctx* ctx, void* host_rgba8, void** out_dev);

/* float4 accumulation buffer {sum_r, sum_g, sum_b, n_samples} per pixel. */
int crt_get_accum(crt_ctx* ctx, float* host_rgba32f, void** out_dev);

/* Overwrite the accumulation buffer (checkpoint resume / after an external reduce). */
int crt_set_accum(crt_ctx* ctx, const float* host_rgba32f);

/* Re-run the reference's writeColor (src/CRTUtility.cuh:21-32) over the accumulation
 * buffer: rgba8 = pack(gamma(sum / n)). Used on rank 0 after the multi-GPU reduce. */
int crt_tonemap(crt_ctx* ctx);

/* replaces CUDARenderer::getRandState (reference: src/CUDARenderer.cuh:18): the RNG state
 * as 6 planes of width*height uint32 (v0..v4, d). host may be NULL. */
int crt_get_rng_state(crt_ctx* ctx, uint32_t* host_6planes, void** out_dev);
int crt_set_rng_state(crt_ctx* ctx, const uint32_t* host_6planes);

int crt_get_stats(crt_ctx* ctx, crt_stats* out);

/* CUDA stream (cudaStream_t) the context launches on, for event timing by the caller. */
int crt_get_stream(crt_ctx* ctx, void** out_stream);

/* Make the context launch on the caller's stream (cudaStream_t) instead of its own, so that the caller's
 * events and collectives (e.g. the multi-GPU reduce of the accumulation buffer) order with the render
 * without extra synchronisation. NULL restores the context's own stream; that also means the legacy default
 * stream (handle 0, what the reference launches on, src/CUDARenderer.cuh:57) cannot be selected: pass a stream
 * the caller created. */
int crt_set_stream(crt_ctx* ctx, void* stream);

/* ---- multi-GPU -------------------------------------------------------------
 * The reference is single-GPU (one device, default stream: src/CUDARenderer.cuh:55-60). Its sample loop
 * (src/CUDAKernels.h:158-162) shards: GPU g of G renders its share of the samples of EVERY pixel from the
 * cuRAND subsequences g*width*height + pixel (crt_init_rng's base), then ONE sum-reduce of the float4
 * accumulation buffers onto a root, which tonemaps. NCCL is bound at run time (libnccl.so.2).
 *
 * (a) one process per GPU (torchrun, mpirun): rank 0 calls crt_comm_unique_id and hands the 128 bytes to
 *     the other ranks; every rank calls crt_comm_init (collective), then per frame
 *     crt_render(spp_g, CRT_RENDER_NO_TONEMAP) and crt_reduce_accum(root) (collective). */
#define CRT_COMM_ID_BYTES 128
int crt_comm_unique_id(void* out_id, size_t bytes);                 /* ncclGetUniqueId */
int crt_comm_init(crt_ctx* ctx, const void* id, int32_t nranks, int32_t rank);   /* ncclCommInitRank on ctx's device */
int crt_comm_destroy(crt_ctx* ctx);
int crt_comm_info(crt_ctx* ctx, int32_t* nranks, int32_t* rank, int32_t* nccl_version);
/* ncclReduce(sum) of the accumulation buffer, in place, on the context's stream (ordered behind the
 * render), onto rank `root`; the root then runs crt_tonemap. Returns after the stream has drained: the
 * buffer may be overwritten by the next frame. Without a communicator it is crt_tonemap. */
int crt_reduce_accum(crt_ctx* ctx, int32_t root);

/* (b) one process, several GPUs: one context per device inside the library, one host thread per device
 *     while rendering. Device devices[0] is the root: after crt_group_render its context
 *     (crt_group_ctx(g, 0)) holds the summed accumulation buffer and the RGBA8 image; read them with the
 *     crt_get_* calls. With peer access the reduce is ONE kernel on the root that reads every shard over
 *     NVLink, adds them in device order and tonemaps in the same pass; otherwise (or with
 *     CRT_GROUP_FORCE_NCCL) ncclReduce + crt_tonemap. */
#define CRT_GROUP_MAX_DEVICES 16
enum { CRT_GROUP_FORCE_NCCL = 1u << 0 };
typedef struct crt_group crt_group;
int  crt_group_create(crt_group** out, const crt_config* cfg /* .device ignored */, const int32_t* devices, int32_t num_devices);
void crt_group_destroy(crt_group* g);
const char* crt_group_last_error(const crt_group* g);
int32_t crt_group_size(const crt_group* g);
crt_ctx* crt_group_ctx(crt_group* g, int32_t index);               /* borrowed */
int crt_group_set_scene(crt_group* g, const crt_scene* scene);       /* uploads + builds the BVH on every device */
int crt_group_set_camera(crt_group* g, const crt_camera* cam);
int crt_group_init_rng(crt_group* g, uint64_t seed);                 /* device i: subsequence base i*width*height */
/* params->spp is the frame's TOTAL sample count: device i renders spp/G (+1 for i < spp%G). */
int crt_group_render(crt_group* g, const crt_render_params* params, uint32_t group_flags);
/* sum of the per-device ray / sample / launch counters (render_ms = slowest device), the device time of the
 * last reduce step, and whether the peer-memory kernel is in use. Any pointer may be NULL. */
int crt_group_get_stats(crt_group* g, crt_stats* out_sum, float* out_reduce_ms, int32_t* out_peer);

/* ---- built-in procedural scenes (BASELINE.json configs; assets/ is absent upstream) ---
 * names: "cornell" (C1), "cornell_blob" (C2/C5), "spheres" (C3), "terrain" (C4).
 * param: scene-specific size knob (0 = default; blob/terrain: grid resolution scale).
 * The returned scene is owned by the library; free with crt_scene_free. */
int crt_scene_builtin(const char* name, int32_t param, crt_scene** out_scene,
                      crt_camera* out_default_camera);
void crt_scene_free(crt_scene* scene);

/* ---- OBJ / MTL ingest -------------------------------------------------
 * replaces SceneManager::loadObject + initMeshes (reference: src/SceneManager.h:100-329),
 * including its material classification (:222-247), the per-file re-normalisation of all
 * meshes loaded so far (:307-325), the material-offset rule (:143-145,177) and the two
 * hard-coded spheres of createRandomWorld (src/CUDAKernels.h:69-73) when
 * add_reference_spheres != 0. */
int crt_scene_load_obj(const char* const* paths, int32_t num_paths,
                       int32_t add_reference_spheres, crt_scene** out_scene,
                       char* err_buf, size_t err_len);

/* ---- reference-exact mode ---------------------------------------------------------------
 * Host-only replay of the PARTITION made by the reference's mesh BVH builder (src/Core/Mesh.cuh:121-219),
 * mesh by mesh (mesh_tri_offset; one mesh when NULL). Any output may be NULL; each has num_triangles entries.
 *   out_order[k]   triangle id at position k of the reference's permuted index array
 *   out_rank[id]   the inverse: the position of triangle id. Among hits with bit-equal t the reference keeps the
 *                  one it tests last (Mesh.cuh:73-74,293) = the larger rank; spheres come after all meshes
 *                  (src/CUDAKernels.h:64-73, Core/BVHNode.cuh:115-156).
 *   out_unseen[id] 1 when the triangle sits in a leaf whose unpadded box (Mesh.cuh:258-264) has zero thickness
 *                  along an axis: AABB::hit rejects tmax <= tmin (Core/AABB.cuh:135-136), the reference never
 *                  tests it.
 * crt_set_scene calls this when crt_config.reference_exact is set. */
int crt_scene_reference_order(const crt_scene* scene, uint32_t* out_order, uint32_t* out_rank, uint8_t* out_unseen);

/* ---- image output (replaces the SFML window) --------------------------
 * Writes the RGBA8 image top-down (i.e. applies the reference's flipVertically,
 * src/WindowManager.h:89) as binary PPM (P6). */
int crt_write_ppm(crt_ctx* ctx, const char* path);
/* Writes sum/n as a little-endian PFM (bottom-up, as PFM specifies). */
int crt_write_pfm(crt_ctx* ctx, const char* path);

/* ---- checkpoint / resume ----------------------------------------------
 * Saves / restores accumulation buffer + RNG state + sample count. */
int crt_checkpoint_save(crt_ctx* ctx, const char* path);
int crt_checkpoint_load(crt_ctx* ctx, const char* path);

#ifdef __cplusplus
}
#endif
#endif /* CRT_API_H */
