/*
 * rt_b200.h -- C ABI of librt_b200.so: the B200 (sm_100a) render path for rust_tracer.
 *
 * This header is the drop-in boundary.  Every entry point replaces one call the
 * reference's host code (src/main.rs, src/gui.rs) makes into its render path; the
 * reference file:line each one stands in for is cited beside it.  Signatures use plain
 * pointers, sizes and POD structs only (no C++ / torch types) so that a Rust `-sys`
 * crate, cgo, JNI or ctypes can bind them directly (INTEGRATION.md shows the Rust
 * binding).
 *
 * Conventions
 *   - every function returns RT_OK (0) or a negative rt_status; rt_last_error() gives
 *     the message of the last failure on the calling thread.  Nothing unwinds across
 *     the ABI (the reference panics instead: src/cli.rs:123, src/math/matrix.rs:106).
 *   - there is NO CPU fallback: without a usable CUDA device every compute entry point
 *     fails with RT_ERR_CUDA.
 *   - calls are synchronous unless the name ends in _device; one handle is used from one
 *     thread at a time (the reference is single-threaded, `Rc<RefCell<..>>` everywhere).
 *   - all matrices are row-major 4x4 f32, column-vector convention
 *     (src/math/matrix.rs:5-9,229-252).
 *   - images are row-major: pixel (u, v) = column u, row v, 3 channels, at
 *     ((v_local * width) + u) * 3 where v_local counts only the rows this call renders
 *     (all rows when world == 1).  The reference's RenderBuffer is column-major
 *     buf[u][v] (src/render.rs:5-19); the host shim transposes.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_B200_ABI_VERSION 4

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = -1,      /* bad argument / malformed scene description        */
    RT_ERR_UNSUPPORTED = -2,  /* e.g. a texture function that is not in rt_colorfn */
    RT_ERR_CUDA = -3,         /* CUDA runtime failure or no device                 */
    RT_ERR_NOMEM = -4,        /* host or device allocation failed                  */
    RT_ERR_INTERNAL = -5,
    RT_ERR_TIMEOUT = -6       /* rt_frame_wait: a peer did not signal in time           */
} rt_status;

/* ---- scene description (mirrors src/scene/{mod,sphere,plane,triangle,cube,material}.rs; copied at rt_scene_create) ---------- */

typedef enum rt_shape_kind {
    RT_SHAPE_SPHERE = 0,   /* src/scene/sphere.rs   : unit sphere under `transform`            */
    RT_SHAPE_PLANE = 1,    /* src/scene/plane.rs    : one-sided infinite plane                 */
    RT_SHAPE_TRIANGLE = 2, /* src/scene/triangle.rs : world-space triangle (transform ignored) */
    RT_SHAPE_CUBE = 3      /* src/scene/cube.rs     : unit cube = 12 fixed triangles           */
} rt_shape_kind;

/*
 * One entry of Scene::shapes (src/scene/mod.rs:22-27,40-44).  The index of a shape in
 * rt_scene_desc.shapes is its id (add_shape assigns id = index) and is the tie-break
 * order of Scene::intersect (src/scene/mod.rs:98-116).
 *
 * `transform` / `inv_transform` are what Sphere/Plane/Cube::set_transform store
 * (src/scene/sphere.rs:100-103, plane.rs:54-57, cube.rs:104-107); the caller computes
 * the inverse (the reference's Gauss-Jordan, src/math/matrix.rs:88-142) so that the
 * device sees bit-identical matrices.  Triangles ignore both (src/scene/triangle.rs:51).
 *
 * p[12]:
 *   plane    : origin(3), normal(3), u axis(3), v axis(3)      (src/scene/plane.rs:8-19,22-42)
 *   triangle : v0(3), v1(3), v2(3), normal(3)                  (src/scene/triangle.rs:7-14,17-39)
 *   sphere / cube : unused
 */
typedef struct rt_shape {
    int32_t kind;      /* rt_shape_kind */
    int32_t material;  /* index into rt_scene_desc.materials */
    float transform[16];
    float inv_transform[16];
    float p[12];
} rt_shape;

/*
 * ColorFun = fn((f32,f32)) -> Color (src/scene/material.rs:3) cannot cross an FFI, so
 * the ABI carries a closed set of device texture functions.
 */
typedef enum rt_colorfn_kind {
    RT_COLOR_CONST = 0,        /* rgb, whatever the texture coordinate                     */
    RT_COLOR_CHECKERBOARD = 1  /* src/my_scene.rs:26-43 : rgb or 0.5*rgb by cell parity    */
} rt_colorfn_kind;

typedef struct rt_colorfn {
    int32_t kind; /* rt_colorfn_kind */
    float rgb[3];
} rt_colorfn;

typedef enum rt_material_kind {
    RT_MATERIAL_PHONG = 0,         /* src/scene/material.rs:24-101  (set_diffuse mutates)   */
    RT_MATERIAL_TEXTURE_PHONG = 1  /* src/scene/material.rs:103-191 (set_diffuse is a no-op) */
} rt_material_kind;

typedef struct rt_material {
    int32_t kind; /* rt_material_kind */
    rt_colorfn ambient;
    rt_colorfn diffuse;
    rt_colorfn specular;
    float power;
    float reflectivity;
    float refraction_index;
} rt_material;

typedef enum rt_light_kind {
    RT_LIGHT_POINT = 0,   /* src/scene/mod.rs:178-221 : casts a shadow ray                */
    RT_LIGHT_AMBIENT = 1  /* src/scene/mod.rs:223-246 : direction (0,0,0), no shadow ray  */
} rt_light_kind;

typedef struct rt_light {
    int32_t kind; /* rt_light_kind */
    float pos[3];
    float color[3];
} rt_light;

typedef struct rt_scene_desc {
    const rt_shape *shapes;
    int32_t n_shapes;
    const rt_material *materials;
    int32_t n_materials;
    const rt_light *lights;
    int32_t n_lights;
    float ambient[3]; /* Scene::ambient (src/scene/mod.rs:50-56) */
} rt_scene_desc;

/* Camera (src/render.rs:155-186).  rt_camera_default fills Camera::new's constants. */
typedef struct rt_camera {
    float origin[3];
    float x_min, x_max, y_min, y_max;
    int32_t x_res, y_res;
} rt_camera;

typedef enum rt_render_flags {
    RT_FLAG_NONE = 0,
    /* test every (ray, shape) pair with the exact arithmetic instead of bound-filtering
     * first; same image by construction, used to audit the filter.                      */
    RT_FLAG_AUDIT_ALL_PAIRS = 1,
    /* evaluate with render_tree.rs's weights instead of render.rs's        */
    RT_FLAG_FOREST_SHADING = 2,
    /* rt_render_device only: the output buffers are WHOLE frames (y_res x x_res x 3); every
     * rendered row is stored at its own place instead of compacted.  The buffer may be a
     * peer-mapped allocation on another GPU (CUDA IPC / peer access): the stores of the last
     * kernel then are the gather of a multi-GPU frame, over NVLink, with no collective.    */
    RT_FLAG_FRAME_ROWS = 4
} rt_render_flags;

typedef struct rt_render_params {
    int32_t depth;     /* --depth (src/cli.rs:55-60); 0 renders black (src/render.rs:43-45) */
    int32_t spp;       /* 1 = the reference's pixel-corner ray; 4 = 2x2 stratified jitter   */
    uint64_t seed;     /* jitter seed, only read when spp == 4                              */
    int32_t rank;      /* interleaved row tiles: this call renders the rows v with          */
    int32_t world;     /*   (v / tile_rows) % world == rank                                 */
    int32_t tile_rows; /* >= 1                                                              */
    int32_t flags;     /* rt_render_flags                                                   */
} rt_render_params;

/* Work counters of one render / forest build (the roofline numerator). */
typedef struct rt_stats {
    uint64_t pixels;        /* pixels rendered by this call (after row tiling)            */
    uint64_t rays;          /* Scene::intersect queries: primary + reflect + refract + shadow */
    uint64_t shadow_rays;   /* of which PointLight::get_energy queries                    */
    uint64_t nodes;         /* ray-tree nodes (rays that hit something, shadows excluded) */
    uint64_t sphere_tests;  /* ray x sphere pairs resolved                                */
    uint64_t plane_tests;
    uint64_t triangle_tests;
    uint64_t cube_tests;    /* each = 12 triangle tests + one ray transform               */
    uint64_t exact_tests;   /* pairs that went through the exact (reference-order) test.  This and the
                               evaluation counters below depend on which hits a sweep found first (distance clip,
                               any-hit shadows) and vary by a fraction of a percent between runs; pixels, rays,
                               shadow_rays and nodes do not                                  */
    double   algorithmic_flops; /* per SURVEY 8d: 54/sphere, 30/triangle, 47|14/plane, 33+12*30/cube, 60+70L/node */
    float    sweep_ms;      /* device time in the nearest-hit sweeps (CUDA events)        */
    float    shade_ms;      /* device time in ray generation, spawn and shading           */
    float    total_ms;      /* device time of the whole call                              */
    uint32_t kernel_launches; /* kernels of this library launched by the call             */
    double   filter_flops;  /* flop the sweeps' bounding-sphere filters executed: 2 x (7 FMA per
                               ray x sphere evaluated, 4 for rays swept against an origin table);
                               spheres = bounds and group spheres, skipped groups not counted */
    uint64_t tested_pairs;  /* ray x shape pairs resolved ONE BY ONE (bounding-sphere filter of that
                               shape's own bound, or exact test); the rest of rays x shapes were
                               dismissed wholesale by a group sphere.  SURVEY 8d credits only these */
    /* The same, split so that a reader can price each pair for what was done to it (SURVEY 8d:
     * "culling must be subtracted"): a pair whose own bounding sphere was evaluated and rejected cost
     * 4 or 7 FMA, not an intersection test; only exact tests are intersections the reference's way. */
    uint64_t bound_evals_sphere;   /* ray x sphere pairs whose own bounding sphere the filter evaluated */
    uint64_t bound_evals_triangle; /* ray x triangle pairs, likewise                                  */
    uint64_t bound_evals_cube;     /* ray x cube pairs, likewise                                      */
    uint64_t hier_evals;           /* ray x (group sphere | quad sphere) evaluations                  */
    uint64_t exact_sphere_tests;   /* exact_tests by kind of shape; the planes' are plane_tests       */
    uint64_t exact_triangle_tests;
    uint64_t exact_cube_tests;
    double   exact_flops;          /* exact tests x their SURVEY 8d price (54 / 30 / 393, planes 47 | 14) */
    double   executed_flops;       /* filter_flops + exact_flops: what the sweeps' FP32 pipe was asked to do */
} rt_stats;

typedef struct rt_scene rt_scene;   /* opaque: device-resident scene tables */
typedef struct rt_forest rt_forest; /* opaque: device-resident ray forest   */

/* ---- library ------------------------------------------------------------------------ */

int rt_abi_version(void);
const char *rt_last_error(void);
/* Number of CUDA devices, or RT_ERR_CUDA. */
int rt_device_count(void);
/* Camera::new(x_res, y_res) (src/render.rs:166-176). */
void rt_camera_default(int32_t x_res, int32_t y_res, rt_camera *out);
/* Rows a (rank, world, tile_rows) call renders out of `height`; fills rows[] (capacity
 * `cap`, may be NULL) in increasing order and returns the count.                         */
int32_t rt_tile_rows_for_rank(int32_t height, int32_t rank, int32_t world, int32_t tile_rows,
                              int32_t *rows, int32_t cap);

/* ---- scene -------------------------------------------------------------------------- */

/* Replaces building `Scene` + `create_scene` for the device (src/main.rs:33-34): copies
 * the description into structure-of-arrays device tables on CUDA device `device`.       */
int rt_scene_create(const rt_scene_desc *desc, int32_t device, rt_scene **out);
/* Material::set_diffuse + friends after upload (src/gui.rs:221-235): replaces the whole
 * material table, geometry untouched.  n must equal the scene's n_materials.  While a forest
 * built from the scene is alive, reflectivity and refraction_index must stay as they are (they
 * decided which rays the forest holds): RT_ERR_UNSUPPORTED otherwise.                   */
int rt_scene_update_materials(rt_scene *scene, const rt_material *materials, int32_t n);
/* Launch this scene's kernels and copies on `stream` (a cudaStream_t created by the caller
 * on the scene's device, e.g. a torch stream's handle) so that the caller's own events
 * bracket the work (a level's shadow sweeps run on side streams of the library, forked from
 * and joined back into `stream` by events inside the call); NULL restores the scene's private
 * stream.  Calls stay synchronous.                                                       */
int rt_scene_set_stream(rt_scene *scene, void *stream);
/* Forests built from the scene keep using its tables: if any is alive, the scene is freed when the
 * last of them is destroyed (either order of the two destroy calls is fine).                  */
void rt_scene_destroy(rt_scene *scene);
/* Destroyed scenes and forests leave their ray pools (device memory) cached, one set per
 * device, for the next scene; rt_trim frees the cache of `device` (-1: every device, the
 * host-side tables kept for the last scene description built, and the idle worker threads
 * kept between rt_multi objects).                                                        */
int rt_trim(int32_t device);

/* ---- Basic renderer ------------------------------------------------------------------ */

/* render::render(&camera, &scene, &mut buffer, depth) (src/render.rs:31-38; callers
 * src/main.rs:251, src/render.rs:27).  Host output buffers, either may be NULL:
 *   rgb_f32 : un-clamped Color per pixel (what RenderBuffer holds), 3 floats
 *   rgb_u8  : Color::as_u8 (src/scene/color.rs:43-46), what bmp.rs writes
 * Both cover only the rows of (rank, world, tile_rows).  stats may be NULL.             */
int rt_render(rt_scene *scene, const rt_camera *camera, const rt_render_params *params,
              float *rgb_f32, uint8_t *rgb_u8, rt_stats *stats);

/* Same, but the outputs are DEVICE pointers on the scene's device (e.g. a torch tensor's
 * data_ptr) and nothing is copied to the host; returns after the work is complete.  Used
 * to time the path with inputs resident, and to hand tiles to an NCCL gather.           */
int rt_render_device(rt_scene *scene, const rt_camera *camera, const rt_render_params *params,
                     void *d_rgb_f32, void *d_rgb_u8, rt_stats *stats);

/* ---- multi-GPU frame ------------------------------------------------------ */

/* One whole frame buffer on the GPU that owns the image (rank 0), mapped into every other
 * rank's process (CUDA IPC, peer access over NVLink): each rank renders its row tiles with
 * rt_render_device + RT_FLAG_FRAME_ROWS straight into it, so the only exchange step of the
 * path -- collecting the tiles for the image writer (the reference has one RenderBuffer,
 * src/render.rs:31-38) -- is done by the stores of the last kernel.
 *   owner:   rt_frame_alloc, rt_frame_export (64-byte handle, sent to the peers by any means)
 *   peers:   rt_frame_import on their own device
 *   both:    rt_frame_release(ptr, imported)
 * Behind the `bytes` of pixels (rounded up to 256) the allocation carries
 * RT_FRAME_SIGNAL_WORDS 32-bit words, zeroed by rt_frame_alloc, through which the ranks tell
 * each other how far they are, on the device, in stream order, with no host barrier and no
 * collective:
 *   rt_frame_signal  enqueues "word := value" after everything already on `stream` (a
 *                    system-scope release: the pixels this rank stored are visible to whoever
 *                    sees the word); returns at once.
 *   rt_frame_wait    enqueues a one-warp kernel that polls words [first_word, first_word +
 *                    n_words) until each is >= value (32-bit wrap-around safe), then waits
 *                    for the stream; RT_ERR_TIMEOUT after timeout_ms (0 or > 60000: 60000),
 *                    so a dead peer cannot hang the GPU.
 * Convention used by rust_tracer_b200/distributed.py:PeerFrame -- word r: rank r has finished
 * frame number `value`; word RT_FRAME_WORD_RELEASED: the owner has consumed it and the
 * peers may overwrite the pixels.  `bytes` must be the value given to rt_frame_alloc.     */
#define RT_FRAME_HANDLE_BYTES 64
#define RT_FRAME_SIGNAL_WORDS 256
#define RT_FRAME_WORD_RELEASED 255
int rt_frame_alloc(int32_t device, size_t bytes, void **d_ptr);
int rt_frame_export(int32_t device, void *d_ptr, uint8_t handle[RT_FRAME_HANDLE_BYTES]);
int rt_frame_import(int32_t device, const uint8_t handle[RT_FRAME_HANDLE_BYTES], void **d_ptr);
int rt_frame_release(int32_t device, void *d_ptr, int32_t imported);
int rt_frame_signal(int32_t device, void *d_ptr, size_t bytes, int32_t word, uint32_t value, void *stream);
int rt_frame_wait(int32_t device, void *d_ptr, size_t bytes, int32_t first_word, int32_t n_words, uint32_t value,
                  uint32_t timeout_ms, void *stream);

/* ---- one host call, several GPUs (SURVEY 8b "Threading", 8e) ---------------------------- */

/* The reference renders a frame with ONE blocking call from one thread, render::render(&camera,
 * &scene, &mut buffer, depth) (src/render.rs:31-38; caller src/main.rs:244-261).  rt_multi is the
 * same call for a host that owns several GPUs: the scene tables are built once on the host and
 * uploaded to every listed device; each render call splits the frame into interleaved tiles of
 * params->tile_rows rows (tile k goes to device k % n: pixels are independent, no collective),
 * drives every device from its own worker thread inside the library and returns when the whole
 * image is in place.  params->rank / world must be 0 / 1.  stats are summed over the devices,
 * the times are those of the slowest one.                                                      */
typedef struct rt_multi rt_multi; /* opaque: one scene resident on n devices */
int rt_multi_create(const rt_scene_desc *desc, const int32_t *devices, int32_t n_devices, rt_multi **out);
int32_t rt_multi_device_count(const rt_multi *multi);
int rt_multi_update_materials(rt_multi *multi, const rt_material *materials, int32_t n);
/* Whole-frame HOST buffers (either may be NULL), y_res x x_res x 3: every device copies its own
 * rows to their place in the image over its own PCIe link, all devices in parallel.          */
int rt_multi_render(rt_multi *multi, const rt_camera *camera, const rt_render_params *params, float *rgb_f32,
                    uint8_t *rgb_u8, rt_stats *stats);
/* Whole-frame DEVICE buffers on the FIRST listed device: every device's last kernel stores its rows
 * there over NVLink / NVSwitch peer access (RT_FLAG_FRAME_ROWS): the gather is those stores.
 * RT_ERR_UNSUPPORTED when a device has no peer access to the first one.                       */
int rt_multi_render_device(rt_multi *multi, const rt_camera *camera, const rt_render_params *params, void *d_rgb_f32,
                           void *d_rgb_u8, rt_stats *stats);
void rt_multi_destroy(rt_multi *multi);
/* rt_multi_create + rt_multi_render + rt_multi_destroy: the reference's call as it stands, a scene
 * description in, an image out (what rust/reference_patch/gpu.rs binds for `--gpus N`).        */
int rt_render_multi(const rt_scene_desc *desc, const int32_t *devices, int32_t n_devices, const rt_camera *camera,
                    const rt_render_params *params, float *rgb_f32, uint8_t *rgb_u8, rt_stats *stats);

/* ---- RayForest renderer -------------------------------------------------------------- */

/* render_tree::generate_ray_forest (src/render_tree.rs:147-164; caller src/main.rs:269):
 * all ray casting; the forest stays on the device.                                      */
int rt_forest_build(rt_scene *scene, const rt_camera *camera, const rt_render_params *params,
                    rt_forest **out, rt_stats *stats);
/* render_tree::render_forest (src/render_tree.rs:121-127; caller src/main.rs:286):
 * shading only, with the scene's CURRENT material table.  `ambient` overrides the
 * scene's ambient when not NULL (the reference passes it explicitly).                   */
int rt_forest_eval(rt_forest *forest, const float *ambient, float *rgb_f32, uint8_t *rgb_u8,
                   rt_stats *stats);
/* render_tree::render_forest_filter (src/render_tree.rs:129-145; callers src/main.rs:187,
 * src/gui.rs:168): re-shades only pixels whose tree holds one of mutated_ids; the other
 * pixels of the in/out buffers are left untouched.  *n_evaluated (may be NULL) receives
 * the number of trees re-shaded.                                                         */
int rt_forest_eval_filtered(rt_forest *forest, const float *ambient, const int32_t *mutated_ids,
                            int32_t n_ids, float *rgb_f32_inout, uint8_t *rgb_u8_inout,
                            uint64_t *n_evaluated, rt_stats *stats);

/* RayForest::stats (src/render_tree.rs:73-105). */
typedef struct rt_forest_stats {
    uint64_t num_trees;
    uint64_t num_intersections;
    uint64_t smallest_tree, largest_tree, median, p90, p95, p99;
} rt_forest_stats;
int rt_forest_get_stats(rt_forest *forest, rt_forest_stats *out);
/* RayForest::size (src/render_tree.rs:62-64) and trees_with (:68-70). */
int rt_forest_size(rt_forest *forest, uint64_t *out);
int rt_forest_trees_with(rt_forest *forest, int32_t shape_id, uint64_t *out);
/* rt_forest_trees_with without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same inputs, same result; the per-tree marks
 * and the count live in a static device array of RT_STATIC_TREES_MAX + 256 bytes that every GPU
 * loading the library holds until exit.  RT_ERR_INVALID for a forest of more than
 * RT_STATIC_TREES_MAX trees.  Calls no cudaMalloc / cudaFree.  All calls on one GPU share the
 * array: calls that overlap on the GPU race, serialise them per device.                 */
#define RT_STATIC_TREES_MAX (4 * 1024 * 1024)
int rt_forest_trees_with_static(rt_forest *forest, int32_t shape_id, uint64_t *out);
void rt_forest_destroy(rt_forest *forest);

/* ---- measurement --------------------------------------------------------------------- */

/* FMA-saturation micro-kernel: measured FP32 throughput of `device` in TFLOP/s (2 flop
 * per FFMA) and the SM count / clock used for the nominal figure.           */
int rt_measure_fp32_peak(int32_t device, double *tflops_measured, int32_t *sm_count,
                         int32_t *sm_clock_khz);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
