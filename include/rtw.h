/*
 * rtw.h — C ABI of the B200 path-tracing backend (drop-in boundary).
 *
 * The reference (Terence-23/RayTracing-in-a-weekend, Rust/ tree) has no FFI: its render path is
 * entered through four Rust functions that take a `Scene`, a `Viewport` and an integrator
 * closure and return `Img = Vec<Vec<Rgb<f32>>>`:
 *     Viewport::render          Rust/src/viewport.rs:430   (stratified, serial)
 *     Viewport::render_no_rand  Rust/src/viewport.rs:479   (1 centre ray / pixel)
 *     async_render              Rust/src/viewport.rs:215   (row tasks, random jitter, motion blur)
 *     render_multi              Rust/src/viewport.rs:249   (frames start_frame..start_frame+n)
 * This header is what a Rust `extern "C"` block (rust_shim/src/ffi.rs), a ctypes loader
 * (raytracing-in-a-weekend_b200/_abi.py) or a C++ caller (include/rtw.hpp) binds instead.
 * Every struct is plain data: counts, floats and pointers to caller-owned arrays. The library
 * deep-copies at rtw_scene_create and owns device memory until rtw_scene_destroy.
 *
 * No entry point falls back to the CPU: when no CUDA device is usable every call returns
 * RTW_E_CUDA and rtw_last_error() says why.
 */
#ifndef RTW_H
#define RTW_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTW_ABI_VERSION 3

/* ---- error codes (negative; 0 = ok). The reference panics/unwraps instead (viewport.rs:242). */
enum {
    RTW_OK = 0,
    RTW_E_INVALID = -1,   /* bad argument / inconsistent scene description */
    RTW_E_CUDA = -2,      /* CUDA runtime error or no device */
    RTW_E_NOMEM = -3,
    RTW_E_UNSUPPORTED = -4
};

/* ---- Material  (Rust/src/objects/materials.rs:15-20; on_hit :105-154) */
typedef struct rtw_material {
    float metallicness;   /* 0 = lambertian, 1 = mirror, between = blend of directions */
    float opacity;        /* > 0 selects the dielectric branch; the value is otherwise unused */
    float ir;             /* refraction index */
    float emitted[3];     /* `emmited` in the reference */
} rtw_material;

/* ---- PerlinNoise tables (Rust/src/texture.rs:63-69). `ranfloat` is dead in the reference. */
typedef struct rtw_noise {
    float    ranvec[256][3];
    uint32_t perm_x[256];
    uint32_t perm_y[256];
    uint32_t perm_z[256];
} rtw_noise;

/* ---- ImageTexture (Rust/src/texture.rs:21-27). `width` is the reference's `row`,
 *      `height` its `col`; texel (x,y) = rgb[(y*width + x)*3 .. +3]  (texture.rs:265). */
typedef struct rtw_texture {
    uint32_t     width;
    uint32_t     height;
    const float* rgb;          /* width*height*3 floats, caller-owned */
    int32_t      noise_id;     /* -1 = None; else index into rtw_scene_desc.noises */
    float        noise_scale;
} rtw_texture;

/* ---- Sphere (Rust/src/objects/sphere.rs:13-20) */
typedef struct rtw_sphere {
    float    origin[3];
    float    radius;
    float    col_mod[3];
    float    velocity[3];
    uint32_t mat_id;
    uint32_t tex_id;
} rtw_sphere;

/* ---- Quad (Rust/src/objects/quad.rs:8-20). normal/d/w are re-derived as quad.rs:93-105.
 *      `velocity` is stored and never read by the reference (quad.rs:37-81); kept for layout parity. */
enum { RTW_SHAPE_PARALLELOGRAM = 0,    /* inside test alfa, beta in [0,1]        quad.rs:57-60 */
       RTW_SHAPE_TRIANGLE      = 1 };  /* inside test alfa, beta >= 0, alfa+beta <= 1
                                        * Rust2/src/objects/triangle.rs:107-113 */
typedef struct rtw_quad {
    float    origin[3];
    float    u[3];
    float    v[3];
    float    velocity[3];
    uint32_t mat_id;
    uint32_t tex_id;
    uint32_t shape;                    /* RTW_SHAPE_*: a triangle is the half of the parallelogram next to
                                        * `origin`; same normal/d/w (triangle.rs:36-47 = quad.rs:93-105),
                                        * same texel rule; its leaf box spans the three corners and is at
                                        * least 0.005 wide (triangle.rs:50-89) */
} rtw_quad;

/* ---- Instance (Rust/src/objects/instance.rs:28-39). An instance owns a range of the scene's
 *      sphere and quad arrays (those ranges lie after the top-level primitives). */
enum { RTW_DENSITY_SURFACE = 0,        /* dist_fn = surface        instance.rs:21-23 */
       RTW_DENSITY_CONST   = 1 };      /* dist_fn = const_density  instance.rs:24-26 */
typedef struct rtw_instance {
    uint32_t first_sphere, n_spheres;
    uint32_t first_quad,   n_quads;
    float    translation[3];
    float    rotation[3];              /* Euler angles fed to Vec3::rotated (vec3.rs:161-181) */
    uint32_t density_kind;
    float    density;
} rtw_instance;

/* ---- one entry of the light list: a top-level sphere or quad of the scene */
enum { RTW_LIGHT_SPHERE = 0, RTW_LIGHT_QUAD = 1 };
typedef struct rtw_light {
    uint32_t kind;                     /* RTW_LIGHT_* */
    uint32_t index;                    /* into spheres[0..n_top_spheres) or quads[0..n_top_quads) */
} rtw_light;

/* ---- Scene (Rust/src/viewport.rs:79-88). spheres[0..n_top_spheres) and quads[0..n_top_quads)
 *      are Scene.spheres / Scene.quads; the rest belong to instances. */
typedef struct rtw_scene_desc {
    uint32_t abi_version;              /* = RTW_ABI_VERSION */
    uint32_t n_spheres, n_top_spheres;
    uint32_t n_quads,   n_top_quads;
    uint32_t n_instances;
    uint32_t n_materials, n_textures, n_noises;
    const rtw_sphere*   spheres;
    const rtw_quad*     quads;
    const rtw_instance* instances;
    const rtw_material* materials;
    const rtw_texture*  textures;
    const rtw_noise*    noises;
    float background[3];               /* Scene.background_color, viewport.rs:87 */
    /* Light list of the light-biased integrators (Rust2/src/viewport/ray_color.rs:55-164 take `lights:
     * Arc<[Arc<dyn Object>]>` from their caller): scene primitives that are ALSO in spheres[0..n_top_spheres) /
     * quads[0..n_top_quads). May be empty (n_lights = 0, lights = NULL). */
    uint32_t n_lights;
    const struct rtw_light* lights;
} rtw_scene_desc;

/* ---- Camera: the derived (private) fields of Viewport (viewport.rs:48-77). Fill by hand or
 *      with rtw_camera_init, which restates Viewport::new (viewport.rs:308-401). */
typedef struct rtw_camera {
    uint32_t width, height;
    float origin[3];
    float pixel00[3];                  /* `upper_left_corner` (a direction; origin NOT added) */
    float delta_u[3], delta_v[3];      /* p_delta_u / p_delta_v */
    float u[3], v[3];                  /* lens basis */
    float lens_radius;
} rtw_camera;

/* ---- integrators: the closures the reference passes as `ray_color` */
enum {
    RTW_INT_GRADIENT   = 0,  /* ray_color_gradient   viewport/ray_color.rs:12-41   mint 1e-3 maxt 1e5 */
    RTW_INT_D          = 1,  /* ray_color_d          main.rs:17-46                 mint 1e-3 maxt 1e3 */
    RTW_INT_BG_COLOR   = 2,  /* ray_color_bg_color   viewport/ray_color.rs:43-92   mint 1e-3 maxt 1e4 */
    RTW_INT_NORMAL     = 3,  /* first-hit normal shading, viewport.rs:524-554      mint 0    maxt 1e3 */
    RTW_INT_ALBEDO     = 4,  /* first-hit col_mod (miss -> background)             mint 1e-3 maxt 1e5 */
    RTW_INT_GLASS_TEST = 5,  /* glass_tests.rs:8-55: non-mirror hit -> yellow, miss -> blue; 1e-3,1e3 */
    /* The two light-biased estimators of the Rust2 tree, over THIS scene model (uber-material, emission
     * in the material): at a hit, one shadow ray per light of rtw_scene_desc.lights towards the middle of the light's
     * box; the light counts when the scene's closest hit along that ray is the light itself, and adds
     * emitted * material_pdf / distance^2. mint 1e-3, maxt 1e3 (Rust2's test scenes, tests.rs:639-644). Megakernel only. */
    RTW_INT_LIGHT_CAST   = 6, /* light_biased_ray_cast  Rust2/src/viewport/ray_color.rs:55-109: first hit only, direct light
                               * averaged over the lights that count, + emitted; miss -> background */
    RTW_INT_LIGHT_BIASED = 7  /* light_biased_ray_color ray_color.rs:111-164: (L(reflected) + sum w * direct) * col_mod /
                               * (1 + n w) + emitted, w = rtw_render_opts.light_weight; depth exhausted or miss -> background */
};

/* ---- which reference render loop to mirror */
enum {
    RTW_MODE_RANDOM     = 0, /* async_render/render_row  viewport.rs:270-305 */
    RTW_MODE_STRATIFIED = 1, /* Viewport::render         viewport.rs:430-478  (time = 0) */
    RTW_MODE_NO_RAND    = 2  /* Viewport::render_no_rand viewport.rs:479-516  (1 ray/pixel) */
};

enum {
    RTW_OUT_RESOLVED = 0,    /* mean over samples, then powf(1/gamma) (viewport.rs:207-213,301) */
    RTW_OUT_SUM      = 1     /* raw per-pixel f32 sums of this call's sample range (multi-GPU) */
};

enum {
    /* Use the kernels compiled with -fmad=false: no a*b+c contraction, i.e. the same f32 operation
     * sequence rustc emits for the reference (parity runs). Default: nvcc's FFMA contraction. */
    RTW_FLAG_STRICT_FP = 1u,
    /* Pipeline choice. Default (neither flag): the wavefront pipeline (extend / shade kernels over
     * path queues in HBM) for renders of at least 2^24 paths of a scene with a tree to walk; the single
     * persistent-lane megakernel below that, and at any size for a scene whose top level is at most eight
     * primitives with no instance level below (measured rule: rtw_api.cu wavefront_pays). The flags
     * force one or the other. */
    RTW_FLAG_WAVEFRONT = 2u,
    RTW_FLAG_MEGAKERNEL = 8u,
    /* Use the instrumented build of the fast kernels, which also counts BVH node visits and
     * primitive tests (rtw_stats.nodes ...): the algorithmic work behind the roofline. Slower. */
    RTW_FLAG_COUNT_WORK = 4u
};

typedef struct rtw_render_opts {
    uint32_t samples;        /* Viewport.samples (total over all devices) */
    uint32_t depth;          /* Viewport.depth */
    float    gamma;          /* Viewport.gamma; resolve applies powf(1/gamma) */
    uint32_t integrator;     /* RTW_INT_* */
    uint32_t mode;           /* RTW_MODE_* */
    float    shutter_speed;  /* Viewport.shutter_speed */
    float    fps;            /* Viewport.fps */
    uint32_t frame;          /* Viewport.frame */
    uint64_t seed;           /* Philox key; the reference is unseeded (thread_rng) */
    float    mint, maxt;     /* used only when maxt > 0; else the integrator's own constants */
    /* sharding: this call renders samples [sample_begin, sample_begin+sample_count)
     * of rows [row_begin, row_begin+row_count). sample_count == 0 means "all"; row_count == 0 means
     * "all rows". The output buffer holds row_count*width*3 floats. */
    uint32_t sample_begin, sample_count;
    uint32_t row_begin, row_count;
    uint32_t output;         /* RTW_OUT_* */
    uint32_t flags;
    float    light_weight;   /* RTW_INT_LIGHT_BIASED: `biased_weight` of light_biased_ray_color (ray_color.rs:116) */
    /* Interleaved row bands in ONE render (tiles across devices): with row_stride != 0 the call renders
     * row_count rows taken in bands of row_band rows every row_stride rows from row_begin on - output row jr is image
     * row row_begin + (jr / row_band) * row_stride + jr % row_band - into a compact row_count*width*3 buffer. Device g
     * of n takes row_begin = g*band, row_band = band, row_stride = n*band. 0 / 0 = one contiguous block of rows. */
    uint32_t row_band, row_stride;
} rtw_render_opts;

typedef struct rtw_stats {
    uint64_t paths;          /* camera samples traced by the last render call */
    uint64_t rays;           /* path segments (Scene::collision_normal equivalents) */
    uint64_t nan_paths;      /* samples whose radiance was non-finite */
    uint64_t gpu_launches;   /* kernels launched by the last render call */
    float    kernel_ms;      /* device time of the trace kernel(s), CUDA events */
    float    total_ms;       /* device time incl. resolve and D2H copy when there is one */
    /* filled only by RTW_FLAG_COUNT_WORK renders, else 0 */
    uint64_t nodes;          /* 4-wide BVH nodes visited (each = four child-box slab tests, 64 B of half-float boxes) */
    uint64_t sphere_tests, quad_tests, instance_tests;
    uint64_t hits;           /* segments that ended on a surface (Hit evaluated + scatter) */
    /* wavefront renders only, else 0 */
    uint64_t iterations;     /* extend + shade rounds */
    float    extend_ms;      /* mean device time of one wf_extend launch (CUDA events, sampled every 8th round) */
    float    shade_ms;       /* mean device time of one wf_shade launch (same samples) */
} rtw_stats;

typedef struct rtw_scene rtw_scene;   /* opaque: device-resident flattened scene + BVH */

/* --- library */
int         rtw_abi_version(void);
const char* rtw_last_error(void);               /* thread-local message of the last failure */
int         rtw_device_count(void);             /* 0 when no CUDA device is visible */

/* --- camera helpers (host only) */
/* Viewport::new (viewport.rs:308-401): aspect_ratio is width/height as f32; height is derived
 * as (width / aspect) truncated. vfov in degrees. */
int rtw_camera_init(rtw_camera* cam, uint32_t width, float aspect_ratio, float vfov_deg,
                    const float origin[3], const float direction[3], const float vup[3],
                    float lens_radius);
/* Viewport::new_from_res (viewport.rs:402-428) */
int rtw_camera_init_res(rtw_camera* cam, uint32_t width, uint32_t height, float vfov_deg,
                        const float origin[3], const float direction[3], const float vup[3],
                        float lens_radius);
void rtw_render_opts_default(rtw_render_opts* o);   /* samples 1, depth 10, gamma 2, fps 30 ... */

/* --- scene lifetime. `device` = CUDA ordinal, or -1 for the calling thread's current device.
 * Replaces Scene::new / new_sphere / new_quad (viewport.rs:89-135): flattens, builds the
 * two-level SAH BVH on the host and uploads. */
int  rtw_scene_create(const rtw_scene_desc* desc, int device, rtw_scene** out);
void rtw_scene_destroy(rtw_scene* scene);
/* Device and pinned memory of destroyed scenes is kept for the next rtw_scene_create / render on the
 * same device (up to RTW_MEM_CACHE_MB, default 4096; 0 disables), so that a scene-per-frame loop
 * (render_multi clones the scene per frame, viewport.rs:263) does not pay cudaMalloc / cudaFree of
 * the path pool every frame. This gives the idle blocks back. `device` < 0 = every device. */
int rtw_trim_memory(int device);

/* Optional on-device BVH rebuild (BASELINE north_star; replaces AABB::new / QuadAABB::new /
 * IAABB::new, objects/aabb/aabb.rs:20-71, qaabb.rs:127-163, iaabb.rs:80-116, for callers that
 * rebuild per frame): every level with two or more leaves is rebuilt on the GPU as an LBVH (Morton
 * sort + Karras tree + treelet restructuring + a 4-wide collapse chosen by dynamic programming; within
 * 2-4 % of the host SAH tree's throughput, 1-2 ms) and walked from now on. Same leaves, same leaf boxes:
 * the closest hit of every ray is unchanged. `out_build_ms` (may be NULL) = device time of the build. */
int rtw_scene_rebuild_lbvh(rtw_scene* scene, float* out_build_ms);
/* Walk the host SAH tree built by rtw_scene_create again. Both calls patch the node arrays and instance levels
 * on the scene's own stream and synchronise it; they are NOT ordered against renders the caller has enqueued
 * on its own streams with rtw_render_device: the scene must be idle (those streams synchronised) when either
 * is called. */
int rtw_scene_use_sah(rtw_scene* scene);

/* --- render. Blocking; `out_rgb` is HOST memory (rows top to bottom, [row][col][rgb], like Img).
 * Replaces Viewport::render / render_no_rand / async_render depending on opts->mode. */
int rtw_render(rtw_scene* scene, const rtw_camera* cam, const rtw_render_opts* opts,
               float* out_rgb);
/* Same, but `d_out_rgb` is DEVICE memory on the scene's device and `stream` a cudaStream_t
 * (0 = default stream). All work is enqueued on `stream` and the call does not synchronise it, so a
 * collective or copy enqueued behind it on the same stream is ordered after the frame. The call is
 * NOT asynchronous for renders that take the wavefront pipeline (see RTW_FLAG_WAVEFRONT): its host loop reads the
 * live-path count of finished iteration batches from pinned memory and keeps launching until the
 * frame has drained, i.e. it blocks the calling thread for almost the whole frame (the megakernel
 * path returns right after its one launch). Stats of this call become valid after the stream syncs.
 * One render per scene at a time: a second call on the same scene must not start before the first
 * has returned. The calling thread's current device is left as it was. */
int rtw_render_device(rtw_scene* scene, const rtw_camera* cam, const rtw_render_opts* opts,
                      float* d_out_rgb, void* stream);
/* render_multi (viewport.rs:249-269): frames first_frame..first_frame+n_frames into
 * out_rgb[n_frames][H][W][3] (host). */
int rtw_render_frames(rtw_scene* scene, const rtw_camera* cam, const rtw_render_opts* opts,
                      uint32_t first_frame, uint32_t n_frames, float* out_rgb);
/* Resolve a (reduced) RTW_OUT_SUM buffer in place on the device: /= total_samples, powf(1/gamma)
 * (render_row's last line, viewport.rs:301). n_pixels = rows*width. */
int rtw_resolve_device(float* d_rgb, uint64_t n_pixels, uint32_t total_samples, float gamma,
                       int device, void* stream);
/* write_img_f32's quantiser (Rust/src/write_img.rs:6-19): (x*255) clamp round, NaN -> 0. Device in, device out. */
int rtw_quantize_device(const float* d_rgb, uint8_t* d_rgb8, uint64_t n_values, int device,
                        void* stream);

/* The same quantiser for a HOST image the caller already holds (what write_img_f32 receives):
 * H2D, quantise on the device, D2H. `device` < 0 = current device. */
int rtw_quantize(const float* rgb, uint8_t* rgb8, uint64_t n_values, int device);
/* rtw_render followed by write_img_f32's quantiser, fused on the device: only the 8-bit image
 * crosses PCIe (a quarter of the bytes). `out_rgb8` is HOST memory, rows*width*3 bytes; the
 * caller PNG-encodes it (write_img.rs:19 `img.save`). opts->output must be RTW_OUT_RESOLVED. */
int rtw_render_rgb8(rtw_scene* scene, const rtw_camera* cam, const rtw_render_opts* opts,
                    uint8_t* out_rgb8);

/* The 8-bit bilateral filter of the Rust2 tree (Rust2/src/postprocessing.rs:70-135
 * `bilateral_filter(img, Proximity::new(size, type))`; ProximityType :11-14). Bytes in, bytes out:
 * height rows of width RGB8 pixels. The window is the reference's: [x-size, x+size) x [y-size, y+size)
 * clipped to the image, a diamond |dx|+|dy| < size inside it for RTW_PROX_EDGES; sigma_spatial =
 * ceil(0.02 * diagonal), sigma_range = mean luminance gradient (a flat image therefore comes out black,
 * as in the reference). Not in place. params (4 floats, may be NULL) receives {spatial, avg_gradient,
 * inv_range, inv_spatial}. */
enum { RTW_PROX_SQUARE = 0, RTW_PROX_EDGES = 1 };
int rtw_bilateral_filter_device(const uint8_t* d_rgb8, uint32_t width, uint32_t height, uint32_t size,
                                int proximity, uint8_t* d_out_rgb8, float* d_params, int device,
                                void* stream);
/* rtw_bilateral_filter_device for callers that may not allocate. NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that
 * forbid cudaMalloc. Same arguments, same bytes out as rtw_bilateral_filter_device, but the filter's scratch
 * (gradient array, weight table, parameters) is a file-scope static device array of csrc/rtw_post.cu instead
 * of stream-ordered allocations: the call makes no cudaMalloc / cudaMallocAsync / cudaFree and only launches
 * kernels on `stream`. Every device that loads the library holds that array (4.3 MiB) until exit. Limits:
 * (width-2)*(height-2) <= RTW_NOALLOC_MAX_INTERIOR_PIXELS (e.g. 1280x720) and size <= RTW_NOALLOC_MAX_WINDOW;
 * beyond either the call returns RTW_E_UNSUPPORTED and rtw_last_error() names the limit. All calls on a device
 * share the array and it keeps the previous call's contents (every value read is rewritten first, nothing has
 * to start at zero): calls whose kernels can overlap on the GPU — different streams without an ordering between
 * them — race; serialise them. */
#define RTW_NOALLOC_MAX_INTERIOR_PIXELS 1046528u   /* 2^20 - 2048 */
#define RTW_NOALLOC_MAX_WINDOW 16u
int rtw_bilateral_filter_device_noalloc(const uint8_t* d_rgb8, uint32_t width, uint32_t height, uint32_t size,
                                        int proximity, uint8_t* d_out_rgb8, float* d_params, int device,
                                        void* stream);
/* The same for HOST images (what the reference function receives and returns): H2D, filter, D2H.
 * out_kernel_ms (may be NULL) receives the device time of the four kernels. `device` < 0 = current. */
int rtw_bilateral_filter(const uint8_t* rgb8, uint32_t width, uint32_t height, uint32_t size,
                         int proximity, uint8_t* out_rgb8, float* out_params, float* out_kernel_ms,
                         int device);

/* ---- several devices behind one call (SURVEY 8b "device list, split {Samples, Tiles}", 8e). The reference's
 * caller is ONE process that calls async_render / render_multi (viewport.rs:215-269, main.rs:76-80,412-414);
 * this is the same call over a list of CUDA devices of that process: the scene is replicated on every
 * device (one SAH build, n uploads), each device renders its share on its own stream from its own host
 * thread, and the shares are combined on devices[0]:
 *   RTW_SPLIT_SAMPLES  device g renders samples [g*spp/n, (g+1)*spp/n) of every pixel into its own f32 sum
 *                      buffer (the Philox counter carries the GLOBAL sample index, so the image does not
 *                      depend on n beyond f32 summation order); one kernel on devices[0] reads the n buffers
 *                      over NVLink peer memory, adds them in device order and resolves (/spp, powf(1/gamma)).
 *   RTW_SPLIT_TILES    bands of `band_rows` rows (0 = 8) dealt round-robin; each device renders its bands as ONE
 *                      render of interleaved rows (rtw_render_opts.row_stride), resolves them and copies them
 *                      straight into their place in out_rgb. No reduction across devices: every
 *                      pixel is summed on one device (equal to a one-device render up to the order of
 *                      that device's own f32 adds).
 *   RTW_SPLIT_FRAMES   (rtw_render_frames_multi) render_multi's frame k -> devices[k mod n]; no exchange.
 * A device may appear more than once in `devices` (its shares then run concurrently on separate streams
 * of that device), which is how a one-GPU box exercises every code path. `out_rgb` is HOST memory. */
typedef struct rtw_scene_multi rtw_scene_multi;
enum { RTW_SPLIT_SAMPLES = 0, RTW_SPLIT_TILES = 1, RTW_SPLIT_FRAMES = 2 };
int  rtw_scene_multi_create(const rtw_scene_desc* desc, const int* devices, int n_devices, rtw_scene_multi** out);
void rtw_scene_multi_destroy(rtw_scene_multi* ms);
int  rtw_render_multi(rtw_scene_multi* ms, const rtw_camera* cam, const rtw_render_opts* opts, uint32_t split,
                      uint32_t band_rows, float* out_rgb);
int  rtw_render_frames_multi(rtw_scene_multi* ms, const rtw_camera* cam, const rtw_render_opts* opts,
                             uint32_t first_frame, uint32_t n_frames, float* out_rgb);
/* Totals over the devices of the last multi render: paths, rays, nan_paths, gpu_launches and the work counters
 * are sums; kernel_ms / total_ms are the maximum over devices (they ran side by side); extend_ms / shade_ms the
 * mean. `per_device` (may be NULL) receives n_devices rtw_stats. combine_ms (may be NULL): device time of the
 * peer-memory combine kernel (samples split), 0 otherwise. */
int  rtw_multi_get_stats(rtw_scene_multi* ms, rtw_stats* total, rtw_stats* per_device, float* combine_ms);
int  rtw_scene_multi_device_count(const rtw_scene_multi* ms);

int rtw_get_stats(rtw_scene* scene, rtw_stats* out);
/* Warp-scheduling counters of the wavefront extend kernel, filled by RTW_FLAG_COUNT_WORK |
 * RTW_FLAG_WAVEFRONT renders: out[0] = loop iterations (per warp, summed); then for the step kinds
 * node, primitive, instance-event, fetch: out[1+k] = times the kind ran, out[5+k] = lanes it ran
 * with in total; out[9+k] = warp cycles spent inside steps of that kind, out[13] = warp cycles in
 * the rest of the loop (ballots, stores). A tuning aid (lanes/steps/32 = SIMD efficiency). */
int rtw_get_sched_stats(rtw_scene* scene, uint64_t out[14]);

/* --- device micro-benchmarks: the denominators of the FP32 and L2 rooflines, which
 * MEASURED_PEAKS.json does not carry. which = RTW_BENCH_FP32 -> TFLOP/s of dependent-free FFMA
 * chains on every SM; RTW_BENCH_L2 -> GB/s of 128-bit read-only loads over a 32 MiB (L2-resident)
 * buffer; RTW_BENCH_HBM -> GB/s of the same loads over 2 GiB. Best of `reps` launches. */
enum { RTW_BENCH_FP32 = 0, RTW_BENCH_L2 = 1, RTW_BENCH_HBM = 2 };
int rtw_microbench(int device, int which, int reps, double* out_value);

/* --- introspection used by tests and the bench (sizes of what was uploaded) */
typedef struct rtw_scene_info {
    uint32_t n_tlas_nodes, n_blas_nodes, n_tlas_leaves, n_flattened_instances;
    uint32_t bvh_max_depth;
    uint32_t n_baked_instances;        /* rigidly placed surface instances whose primitives the fast kernels walk as
                                        * world-space copies in the top level (same closest hit; rtw_host.h) */
    uint64_t device_bytes;             /* device memory held by the scene: scene arrays, textures and (after the first
                                        * wavefront render) the path pool */
    uint64_t h2d_bytes;                /* bytes rtw_scene_create copied host -> device (scene arrays + texels) */
} rtw_scene_info;
int rtw_scene_get_info(rtw_scene* scene, rtw_scene_info* out);

#ifdef __cplusplus
}
#endif
#endif /* RTW_H */
