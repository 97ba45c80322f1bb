/* rt_b200.h — C ABI of the B200-native per-pixel render path.
 *
 * Drop-in boundary: the reference has no FFI layer; the seam this library replaces is the
 * C++ call `tracer.render(framebuffer)` in raytracer/src/main.cpp:74-82 (Tracer::render,
 * raytracer/src/tracer.cpp:247-305) together with the loader call that builds its inputs
 * (jsonio::load_scene_from_json, raytracer/src/json_loader.cpp:499-504).  INTEGRATION.md shows
 * the ~20-line change to the reference's main.cpp that binds these entry points.
 *
 * Conventions: every function returns RT_OK (0) or a negative rt_status; the message of the
 * last failure on the calling thread is rt_last_error().  No C++ exception crosses this ABI.
 * All pointers are plain host pointers unless the name says `_device`.  The caller owns every
 * host buffer; device memory is owned by the library.  There is no CPU fallback: without a
 * CUDA device every compute entry point fails with RT_ERR_CUDA.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_ARG = -1,      /* null/invalid argument */
    RT_ERR_IO = -2,       /* file could not be opened / written           (main.cpp exit code 4 for PNG) */
    RT_ERR_PARSE = -3,    /* JSON parse / processing error                 (main.cpp exit code 3)       */
    RT_ERR_LIMIT = -4,    /* scene exceeds a compiled-in device limit (CSG nesting, recursion depth)   */
    RT_ERR_CUDA = -5,     /* CUDA runtime failure or no device                                           */
    RT_ERR_STATE = -6     /* call order (e.g. render before init)                                        */
} rt_status;

/* RenderMode (raytracer/src/tracer.h:8-11) */
enum { RT_MODE_STANDARD = 0, RT_MODE_PAPER = 1 };

/* ---------------------------------------------------------------------------------------------
 * Flat scene description = output of the host scene compiler, input of rt_scene_upload.
 * Everything is plain-old-data; doubles are kept only where the primary-ray set-up needs them.
 * ------------------------------------------------------------------------------------------- */

/* Material (raytracer/src/geometry.h:5-22) after json_loader.cpp:83-123; `ambient` is already
 * multiplied by the scene ambient (shading.cpp:39).  kd is always 1 in loaded scenes.          */
typedef struct rt_material {
    float albedo[3]; float ks;
    float ambient[3]; float kr;
    float kt; float shininess; float ior; float kd;
} rt_material;

/* PointLight (raytracer/src/scene.h:22-27) */
typedef struct rt_light { float pos[3]; float pad0; float intensity[3]; float pad1; } rt_light;

/* DirectionalLight (raytracer/src/scene.h:10-19) as shade_lambert_phong uses it (shading.cpp:46-76):
 * to_light = (-dir).normalized() (core.h:95-101: length <= 1e-6 falls back to +Y), radiance without falloff.
 * The reference's JSON loader never creates one (json_loader.cpp:191-204); a C++ host that fills
 * scene.dir_lights itself binds rt_compiled_add_dir_light.                                              */
typedef struct rt_dir_light { float to_light[3]; float pad0; float radiance[3]; float pad1; } rt_dir_light;

/* Primitive kinds and flags */
enum { RT_PRIM_SPHERE = 0, RT_PRIM_HALFSPACE = 1, RT_PRIM_POKEBALL = 2, RT_PRIM_PROGRAM = 3 };
enum {
    RT_FLAG_XFORMED = 1,  /* at least one transform node above: front_face is re-derived from the oriented
                             normal (transform.cpp:41,63-64), half-space +-inf endpoints collapse to t=0      */
    RT_FLAG_AFFINE  = 2,  /* sphere/pokeball under a non-uniform scaling: solved in its local frame via `xf`  */
    RT_FLAG_WRAPPED = 4,  /* top-level object is a transform node: child queried on [0,inf], then the hit is
                             accepted only for tmin < t < tmax (transform.cpp:29,38)                           */
    RT_FLAG_NEVER   = 8   /* degenerate scaling (|s| < 1e-6) above: never hit (transform.cpp:103)              */
};

/* One primitive record.  Used both for top-level objects (scene.objects order = JSON order) and
 * for the leaves of CSG programs.
 *   sphere / pokeball : geom = cx, cy, cz, r          (Sphere, geometry.h:79-108)
 *   half-space        : geom = nx, ny, nz, n.p0       (HalfSpace, geometry.h:111-158; n unit)
 *   program           : geom = conservative bounding sphere (cx, cy, cz, R) of the program's leaves, or
 *                       R < 0 when there is none (a half-space or a non-uniformly scaled leaf inside);
 *                       aux = first op, xf = op count                                            */
typedef struct rt_prim {
    float   geom[4];
    int32_t kind;     /* RT_PRIM_*                                                     */
    int32_t flags;    /* RT_FLAG_*                                                     */
    int32_t mat;      /* material id; pokeball: first of 5 (top,bottom,belt,ring,button) */
    int32_t aux;      /* pokeball: index into pokeballs[]; program: first op           */
    int32_t xf;       /* RT_FLAG_AFFINE: index into affines[]; program: op count; else -1 */
    int32_t pad[3];
} rt_prim;

/* Pokeball region parameters (geometry.h:161-242, geometry.cpp:163-180) with the acos removed:
 * ang <= btnOuter  <=>  u.btnDir >= cos_outer ;  ang >= inner  <=>  u.btnDir <= cos_inner.        */
typedef struct rt_pokeball {
    float btn_dir[3]; float cos_outer;
    float up[3];      float cos_inner;   /* `up` is the local +Y axis carried through baked rotations */
    float belt_half;  float pad[3];
} rt_pokeball;

/* world -> leaf-local affine map (rows of a 3x4 matrix), for RT_FLAG_AFFINE leaves */
typedef struct rt_affine { float m[3][4]; } rt_affine;

/* CSG program = postfix evaluation of a csg / transform subtree (csg.cpp:61-163):
 *   RT_OP_LEAF          acc = interval(leaf arg)
 *   RT_OP_COMBINE_LEAF  acc = acc (op) interval(leaf arg)
 *   RT_OP_PUSH          push acc
 *   RT_OP_COMBINE_POP   acc = pop (op) acc
 *   RT_OP_XFIX          a transform node sat here: non-finite endpoints -> 0 (transform.cpp:79-80);
 *                       arg != 0xFFFFF: zero normals become fallbacks[arg] (core.h:58-64 via
 *                       transform.cpp:117-118,201-203)
   RT_OP_COMBINE_SPHERES  a run of COMBINE_LEAF over consecutive plain-sphere leaves, same csg op:
 *                       arg = first_leaf | count << 14   (first_leaf < 16384, count < 1024)
 * word = opcode | csg_op << 4 | arg << 8                                                           */
enum { RT_OP_LEAF = 0, RT_OP_COMBINE_LEAF = 1, RT_OP_PUSH = 2, RT_OP_COMBINE_POP = 3, RT_OP_XFIX = 4, RT_OP_COMBINE_SPHERES = 5 };
enum { RT_CSG_UNION = 0, RT_CSG_INTERSECTION = 1, RT_CSG_DIFFERENCE = 2 };
#define RT_OP_WORD(opcode, csg, arg) ((uint32_t)(opcode) | ((uint32_t)(csg) << 4) | ((uint32_t)(arg) << 8))
#define RT_MAX_CSG_STACK 8      /* right-nesting depth of a CSG program                   */
#define RT_MAX_RECURSION 16     /* medium.recursion above this is rejected (RT_ERR_LIMIT) */

typedef struct rt_scene_desc {
    /* camera + screen (camera.h:8-36); kept in double so primary rays are rounded once */
    double eye[3]; double screen_p[3]; double lx, ly;
    int32_t width, height;                       /* ScreenSpec::nx(), ny() (camera.h:19-21) */
    /* medium (scene.h:40-47) */
    float background[3]; float medium_index;
    int32_t recursion;
    int32_t n_objects, n_leaves, n_pokeballs, n_affines, n_ops, n_materials, n_lights, n_fallbacks;
    const rt_prim*     objects;     /* top-level, JSON order */
    const rt_prim*     leaves;      /* leaves of CSG programs */
    const rt_pokeball* pokeballs;
    const rt_affine*   affines;
    const uint32_t*    ops;
    const rt_material* materials;
    const rt_light*    lights;
    const float*       fallbacks;   /* n_fallbacks x 4 floats (unit normal, pad) */
    /* directional lights (scene.h:37): shaded BEFORE the point lights, in this order (shading.cpp:46,79) */
    int32_t n_dir_lights; int32_t pad0;
    const rt_dir_light* dir_lights;
} rt_scene_desc;

typedef struct rt_scene rt_scene;          /* a scene resident on the device(s) */
typedef struct rt_compiled rt_compiled;    /* host-side compiled scene (owns an rt_scene_desc) */

/* Per-render statistics (device counters + CUDA-event timing). */
typedef struct rt_stats {
    double kernel_ms;          /* trace (+ post-process) kernels, CUDA events on the render stream */
    double total_ms;           /* whole rt_render call incl. copies, host clock                     */
    uint64_t primary_rays;     /* Scene::intersect-equivalent queries for camera rays               */
    uint64_t shadow_rays;      /* Scene::occluded-equivalent queries                                 */
    uint64_t secondary_rays;   /* reflection + refraction closest-hit queries                        */
    uint64_t prim_tests;       /* primitive tests executed (sphere/plane/pokeball/leaf)              */
    int32_t width, height, samples_per_pixel, n_tiles;
} rt_stats;

/* ---- library / device ---------------------------------------------------------------------- */
/* Select the CUDA device this process renders on (one process per GPU).  device < 0: current. */
int rt_init(int device);
/* Contexts on several devices of one box for rt_render_multi (devices = NULL: 0 .. n_devices - 1); the first becomes the
 * process's render device when rt_init has not chosen one.  The reference is single-threaded CPU code (main.cpp:42-94):
 * nothing to replace; SURVEY.md section 5 asks for the GPU list to come from the environment (RAY_GPUS, csrc/ray_main.cpp). */
int rt_init_devices(const int* devices, int n_devices);
/* Releases every device resource of the library; scenes uploaded before become invalid (rt_scene_free stays legal). */
void rt_shutdown(void);
const char* rt_last_error(void);
const char* rt_version(void);

/* ---- host scene compiler (replaces jsonio::load_scene_from_json[_text], json_loader.cpp:458-504) */
int rt_compile_json_text(const char* json_text, rt_compiled** out);
int rt_compile_json_file(const char* path, rt_compiled** out);
const rt_scene_desc* rt_compiled_desc(const rt_compiled*);
/* scene.dir_lights.push_back({dir, radiance}) (scene.h:10-19,37) on a compiled scene; `dir` is the direction the
 * light travels (from the light toward the scene), as in the reference.  Call before rt_scene_upload.   */
int rt_compiled_add_dir_light(rt_compiled*, const double dir[3], const double radiance[3]);
void rt_compiled_free(rt_compiled*);

/* ---- device scene ---------------------------------------------------------------------------- */
int rt_scene_upload(const rt_scene_desc* desc, rt_scene** out);
void rt_scene_free(rt_scene*);
int rt_scene_size(const rt_scene*, int* width, int* height);

/* ---- render (replaces Tracer::render, tracer.cpp:247-305) ------------------------------------
 * Tiles are runs of RT_TILE_PIXELS consecutive pixels in the reference's loop order
 * (standard: y outer from the BOTTOM row, x inner; paper: rows from the top); a rank renders
 * tiles rank, rank+world, ... (interleaved).  world = 1 renders the whole frame.
 *
 * rt_render: host outputs, any of which may be NULL.
 *   rgb8   : H*W*3 bytes, row 0 = top, R,G,B = toByte(channel) (core.h:313-316; main.cpp:19-34 minus
 *            OpenCV's BGR swap)
 *   rgbf   : H*W*3 floats, the linear framebuffer Tracer::render fills
 *   ids    : H*W*S int32 primary-hit codes (object*8 + pokeball region, -1 miss), S = 8 / 1
 *   With world > 1 only this rank's pixels are written (others left untouched).                  */
#define RT_TILE_PIXELS 156
int rt_render(rt_scene*, int mode, int rank, int world,
              uint8_t* rgb8, float* rgbf, int32_t* ids, rt_stats* stats);

/* Host-direct output.  When `rgb8` of rt_render is pinned host memory that is mapped into the device's address space
 * (rt_host_frame_alloc / rt_host_frame_open below, cudaHostAlloc, cudaHostRegister(..Mapped)) the render kernel stores the
 * packed R,G,B bytes straight into it, tile by tile as aligned 128-byte PCIe writes, while the other tiles are still being
 * traced: no device frame, no pack pass, no copy afterwards — this is framebuffer_to_mat_bgr8 (main.cpp:19-34) fused into
 * the kernel's own store.  With world > 1 the pixels of other ranks are left untouched, so several GPUs — one process each
 * (same shm_name) or all in one process (rt_render_multi) — fill ONE frame, each over its own PCIe link.
 *   rt_host_frame_alloc(bytes, NULL, &p)      private pinned + mapped host memory
 *   rt_host_frame_alloc(bytes, "/name", &p)   the same, backed by POSIX shared memory that other processes open with
 *   rt_host_frame_open("/name", bytes, &p)    (every process registers the mapping with its own CUDA context)
 *   rt_host_frame_free(p)                     waits for the devices, unmaps; the creator also unlinks the name            */
int rt_host_frame_alloc(size_t bytes, const char* shm_name, void** host_ptr);
int rt_host_frame_open(const char* shm_name, size_t bytes, void** host_ptr);
int rt_host_frame_free(void* host_ptr);

/* Tracer::render (tracer.cpp:247-305) on n_devices GPUs of one box from ONE process: device i renders tiles i, i + n, ...
 * of a replica of the scene and stores its pixels into `rgb8` (H*W*3, row 0 = top) — directly when rgb8 is a host frame,
 * else through a pinned frame of the library that is copied out at the end.  No peer access, no collective.
 * stats (optional): counters summed over the devices, kernel_ms = the slowest device.                                 */
int rt_render_multi(const rt_scene_desc* desc, int mode, const int* devices, int n_devices, uint8_t* rgb8, rt_stats* stats);

/* Same, but the frame stays on the device: rgba8_device is H*W uchar4 (A = 255), rgbf_device is
 * H*W float4 (may be NULL), ids_device H*W*S int32 (may be NULL).  `stream` is a cudaStream_t
 * (NULL = the library's stream).  Asynchronous unless stats is given.  ONE frame is in flight per device: the tile
 * counter, the ray counters, the scratch buffers and the constant scene header are per device, so a call on another
 * stream first waits (cudaStreamWaitEvent) for the previous frame of that device; calls are serialised by a mutex. */
int rt_render_device(rt_scene*, int mode, int rank, int world,
                     void* rgba8_device, void* rgbf_device, void* ids_device,
                     void* stream, rt_stats* stats);

/* Paper-mode frame (RT_MODE_PAPER) for callers that cannot allocate: code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc.  This entry point has NO counterpart in the reference.
 * Same pixels as rt_render_device(scene, RT_MODE_PAPER, rank, world, rgba8_device,
 * NULL, NULL, stream, NULL), but the G-buffer between the two passes lives in a file-scope static __device__ array
 * instead of a grown cudaMalloc buffer, and the call itself performs no cudaMalloc / cudaFree / allocating library call.
 *   - limit: width * height <= RT_NOALLOC_MAX_PIXELS (1280 x 1024; 20 B of G-buffer per pixel + 256 B of counters =
 *     25 MiB of device memory held by every GPU that loads the library, until the process exits); larger frames are
 *     rejected with RT_ERR_LIMIT and a message naming the limit;
 *   - rt_init and rt_scene_upload (which do allocate) must have been called before; rgba8_device is the caller's
 *     H*W uchar4 buffer; stream as in rt_render_device; asynchronous: order your reads after `stream`;
 *   - every call on a device shares the one scratch array (and, like every render call, the constant scene header):
 *     calls that overlap on the GPU, on different streams, race.  The counters are cleared by every call; the
 *     G-buffer keeps the previous call's contents and needs no clearing (pass 1 writes every entry pass 2 reads). */
#define RT_NOALLOC_MAX_PIXELS (1280 * 1024)
int rt_render_paper_device_noalloc(rt_scene*, int rank, int world, void* rgba8_device, void* stream);

/* ---- peer frame (multi-GPU, one process per GPU) ---------------------------------------------------
 * The reference writes one std::vector<Color> (tracer.cpp:250,297); with several GPUs the ranks write their
 * tiles straight into ONE RGBA8 frame that lives on rank 0, over NVLink peer memory, so the render kernel's
 * pixel store is also the "gather" and no collective moves pixels afterwards:
 *   rank 0      : rt_peer_frame_create(bytes, &ptr, handle)   (zero-filled device allocation + 64-byte IPC handle)
 *   other ranks : rt_peer_frame_open(handle, &ptr)            (maps rank 0's allocation, enables peer access)
 *   every rank  : rt_render_device(scene, mode, rank, world, ptr, ...), then any stream-ordered barrier
 *   at the end  : rt_peer_frame_close(ptr, is_owner)
 * The handle is plain bytes and travels over whatever channel the host has (torch.distributed here).     */
#define RT_IPC_HANDLE_BYTES 64
int rt_peer_frame_create(size_t bytes, void** device_ptr, unsigned char handle[RT_IPC_HANDLE_BYTES]);
int rt_peer_frame_open(const unsigned char handle[RT_IPC_HANDLE_BYTES], void** device_ptr);
int rt_peer_frame_close(void* device_ptr, int is_owner);

/* Known-answer entry point: trace n arbitrary rays (6 doubles each: origin, direction) through
 * Scene::intersect(1e-4, inf) + Tracer::trace; out = 16 doubles per ray:
 *   hit, t, p[3], n[3], front_face, code, r, g, b, occluded on (1e-3, 5], 2 pad.                  */
int rt_trace_rays(rt_scene*, int n, const double* rays, double* out);

/* First n jitter pairs of the Standard loop starting at sample `first_sample` (tracer.cpp:284-293),
 * produced by the DEVICE generator; out = 2n doubles (dx, dy).                                   */
int rt_debug_jitter(uint64_t first_sample, int n, double* out);

/* HOST restatement of the block generator used by the kernels (same inline code, csrc/mt19937.cuh):
 * the 624 tempered outputs of block `block` (outputs 624*block .. +623) of mt19937(seed).  No GPU needed. */
int rt_debug_mt_block_host(uint32_t seed, uint64_t block, uint32_t* out624);

/* HOST restatement of the jump-ahead the checkpoint table is built with (csrc/mt19937.cuh jump_word, polynomials from
 * csrc/mt_jump_gen.cpp): the state 2^log2_twists twists after the seeded state is obtained in ONE step, then twisted once
 * and tempered, i.e. the result must equal rt_debug_mt_block_host(seed, 2^log2_twists, .).  No GPU needed.            */
int rt_debug_mt_jump_host(uint32_t seed, int log2_twists, uint32_t* out624);

/* HOST-ONLY view of the device layout rt_scene_upload builds (no GPU needed).  Long runs of plain spheres /
 * pokeballs are reordered spatially and get one bounding sphere per 16 primitives (the reference scans linearly,
 * scene.cpp:10-42; results are identical).  order[i] = JSON index of device object i; runs4 = (kind,
 * first, count, 1 + first bound or 0) per run; clusters4 = (cx, cy, cz, R): one bound per 16 primitives of a clustered
 * run, one per object of a long run of generic objects (at most n_objects + n_objects/16 + 1 in all).  Outputs may be NULL. */
int rt_debug_device_layout(const rt_scene_desc* desc, int* order, int* n_runs, int* runs4, int* n_clusters, float* clusters4);

/* HOST-ONLY: the table the device's conservative candidate scan reads (csrc/rt_device.cuh sphere_mask): for every entry of
 * the object, CSG-leaf and bound arrays, in that order, (cx, cy, cz, |c|^2 - r^2 - 2^-16 (|c|^2 + r^2)): the second term is
 * the primitive's share of the scan's rounding margin; *scale = max(|c|^2 + r^2) over the finite sphere-like entries.  The scan only selects candidates for the exact
 * arithmetic of Sphere::intersect (geometry.cpp:12-37).  masks4 holds *n x 4 floats (n = objects + leaves + bounds);
 * pass masks4 = NULL to query n.                                                                              */
int rt_debug_mask_table(const rt_scene_desc* desc, int* n, float* masks4, float* scale);

/* Test hooks.  name: "flat", "cull", "wavefront", "global_scene", "grid" (kernel / data-path selection of scenes uploaded
 * afterwards; "cull" = 0 switches the round-level culling off: every ray scans every run; "grid" = 0 / 1: without / with the
 * uniform grid for incoherent rays wherever one can be built) or "mt_sequential" (checkpoint table by the twist-by-twist walk); value -1 = the library's own choice,
 * 0 / 1 = forced off / on.  The library never reads its behaviour from the environment.                          */
int rt_debug_set_option(const char* name, int value);

/* Measured FP32 FFMA peak of the current device in TFLOP/s (dependent-free FFMA micro-benchmark). */
int rt_measure_fp32_peak(double* tflops);

/* ---- PNG (replaces cv::imwrite in main.cpp:84-89) ------------------------------------------- */
int rt_write_png_rgb8(const char* path, const uint8_t* rgb8, int width, int height);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
