/*
 * rt_b200.h — C ABI of the B200-native ray-scene kernel.
 *
 * Drop-in boundary for ONE hot path of blubs/OpenGLRaytracer: the per-pixel
 * compute shader raytrace_compute.glsl and the five host lines that launch it
 * (main.cpp:220-238: glUseProgram / glBindImageTexture(RGBA8) / glUniform1f(time)
 * / glDispatchCompute(W,H,1) / glFinish).  The reference has no FFI; the structs
 * below mirror the shader's own structs field-for-field so that a maintainer can
 * replace the GL dispatch by one call.
 *
 * Citation shorthand: glsl:N = OpenGLRaytracer/raytrace_compute.glsl line N,
 * main.cpp:N = OpenGLRaytracer/main.cpp line N.
 *
 * All structs are plain-old-data, tightly packed 4-byte words.  No torch types,
 * no C++ types.  Every entry point returns RT_OK (0) or a negative rt_status;
 * nothing ever calls exit() (the reference's only error path, main.cpp:50-53).
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 1

/* Limits.  The reference's stack machine holds 100 frames (glsl:873-874) and its
 * runaway guard trips after 10001 loop iterations (glsl:1076-1101); we accept
 * depth 0..RT_MAX_DEPTH and reproduce the in-band red pixel when the guard would
 * have tripped.  */
#define RT_MAX_DEPTH   15
#define RT_MAX_LIGHTS  64

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_BAD_ARG = -1,      /* null pointer, non-positive size, depth out of range ... */
    RT_ERR_CUDA = -2,         /* a CUDA runtime call failed; see rt_last_error_string() */
    RT_ERR_NO_DEVICE = -3,    /* no CUDA device / kernel image not loadable (no CPU fallback) */
    RT_ERR_NOMEM = -4,
    RT_ERR_UNSUPPORTED = -5
} rt_status;

/* glsl:56-69 `struct Material` (20 words). */
typedef struct rt_material {
    float ambient[4];
    float diffuse[4];
    float specular[4];
    float shininess;
    float emissive[4];
    float reflectivity;
    float transparency;
    float refraction_index;
} rt_material;

/* glsl:244-258 `struct Object` with glsl:166-175 Box/Sphere inlined (33 words).
 * Null conventions as in the shader (glsl:178-179, 749, 759): an object is a box
 * iff (box_mins,box_maxs) != all-zero; else a sphere iff sphere_radius != -1;
 * else it is skipped. */
typedef struct rt_object {
    float box_mins[3];
    float box_maxs[3];
    float sphere_radius;
    float position[3];
    float angles[3];          /* pitch, yaw, roll in degrees (glsl:492-503) */
    rt_material material;
} rt_object;

/* glsl:190-196 `struct Light` (15 words). */
typedef struct rt_light {
    float position[3];
    float ambient[4];
    float diffuse[4];
    float specular[4];
} rt_light;

/* The shader's global scene: objects[] + objects_count (glsl:261-321),
 * lights[] + lights_count (glsl:199-224), MAX_RAYTRACE_DEPTH (glsl:22). */
typedef struct rt_scene {
    const rt_object* objects;
    int32_t objects_count;
    const rt_light* lights;
    int32_t lights_count;
    int32_t max_depth;
} rt_scene;

/* glsl:36-50 `struct Camera` (10 words); the "camera uniforms". */
typedef struct rt_camera {
    float position[3];
    float angles[3];          /* pitch, yaw, roll in degrees */
    float v_fov;
    float aspect;
    float near_plane;
    float far_plane;
} rt_camera;

/* Counters filled when rt_options.stats != NULL (instrumented twin kernel). A
 * "ray" is one get_closest_collision() invocation of the reference (glsl:738):
 * primary + reflection + refraction + shadow rays. */
typedef struct rt_stats {
    uint64_t rays;            /* get_closest_collision calls the reference would make */
    uint64_t nodes;           /* bounce-tree nodes (primary+reflection+refraction rays) */
    uint64_t hit_nodes;       /* nodes whose ray hit something */
    uint64_t shadow_rays;
    uint64_t box_tests;       /* ray/box slab tests actually executed on the device */
    uint64_t sphere_tests;    /* ray/sphere tests actually executed on the device */
    uint64_t red_pixels;      /* samples that tripped the runaway guard (glsl:1098) */
    uint64_t rays_not_traversed; /* shadow rays of `rays` whose outcome cannot change the pixel (zero-weight
                                 surface, light without diffuse/specular) and that were not traversed */
    float kernel_ms;          /* CUDA-event time of the render kernel(s) of this call */
    float total_ms;           /* CUDA-event time incl. uploads and the D2H copy */
    uint64_t cull_tests;      /* bounding-sphere / discriminant tests executed for live rays (large scenes) */
    uint64_t warp_iters;      /* bounce-loop iterations summed over warps; nodes/(32*warp_iters) = lane use */
    uint64_t packet_tests;    /* (object, cone) tests of the packet loops, one object per lane (large scenes) */
    uint64_t group_passes;    /* bounce rays: passes over the scene made for a group of rays (per warp) */
    uint64_t loop_passes;     /* bounce rays: passes made by the per-ray loops for what the groups left over (per warp) */
    uint64_t candidates;      /* objects that passed a packet test and went through the per-ray tests (per warp) */
    uint64_t shadow_group_passes;      /* bounce rays: passes made for a group of neighbouring hit points' shadow rays (per warp) */
    uint64_t shaded_lights;   /* (hit node, light) pairs whose diffuse + specular terms were evaluated (glsl:822-833): the
                                 unoccluded ones among the traversed shadow rays (shadow_rays - rays_not_traversed) */
} rt_stats;

/* Extended options; zero-initialise then override (rt_options_default()). The
 * defaults reproduce the shader exactly: 1 sample at the integer pixel corner
 * (glsl:377-378), whole image, host output, synchronous. */
typedef struct rt_options {
    int32_t spp;              /* samples per pixel; 1 = reference behaviour; n*n stratified jitter otherwise */
    uint32_t seed;            /* jitter hash seed (ignored when spp == 1) */
    int32_t y0, y1;           /* row band [y0,y1) to render; y1 <= 0 means height */
    int32_t out_is_device;    /* out_rgba is a device pointer on `device` (no D2H copy) */
    int32_t out_band_only;    /* out_rgba holds only rows [y0,y1) (band-local), else the full image */
    int32_t device;           /* CUDA device ordinal; -1 = current device */
    int32_t strict;           /* 1 = strict-arithmetic kernel (-fmad=false, IEEE divides, shader op order) */
    int32_t async;            /* 1 = do not synchronise (device output only); pair with rt_synchronize() */
    void* stream;             /* cudaStream_t to launch on; NULL = the context's own stream.  Scene uploads run on the
                               * context's stream; a render on another stream is ordered after the last upload, and the
                               * next upload after that render, by events (one caller stream at a time per context) */
    rt_stats* stats;          /* optional counters/timings */
    int32_t host_store;       /* page-locked host out_rgba: RT_HOST_STORE_AUTO (0) lets the kernel store the frame
                               * straight into it over PCIe when the frame is heavy enough to hide the traffic,
                               * _ALWAYS (1), _NEVER (2: device frame + device->host copy) */
    int32_t no_cull;          /* A/B switches, same image: bit 0 = large scenes run every slab test (no bounding-sphere
                               * test first); bit 1 = shadow rays do not use the per-light records; bit 2 = no packet
                               * (cone) tests, every ray meets every bounding sphere itself; bit 3 = no CTA-level packet
                               * tests (each warp scans the scene for its own packets); bit 4 = CTA-level packet tests
                               * also below RT_BLOCK_MIN_OBJECTS objects (test switch); bit 5 = shadow packets without
                               * their fan planes (cone and reach only); bit 6 = bounce rays are not grouped into
                               * packets (per-ray loops); bit 7 = small scenes: shadow rays that start inside a first box
                               * that holds every lit light (the room) are tested against it all the same; bit 8 = the
                               * packet tests go object by object, not through the bounding spheres of groups of 32
                               * neighbouring objects first (built from 2048 objects up, rt_groups.cu); bit 9 = block
                               * kernel: one shadow cone per warp and light, also for a warp whose hit points lie at
                               * several depths (otherwise up to twelve cones, one per cluster of hit points) */
    void* done_flag;          /* optional device pointer to a 32-bit completion word -- typically in ANOTHER GPU's memory,
                               * mapped with rt_ipc_import: done_value is stored there (system-scope release, a one-thread
                               * kernel in stream order behind the render) once every framebuffer store of the call is
                               * visible; pair with rt_wait_flags() */
    uint32_t done_value;
} rt_options;

/* Environment variables (read once per process; profiling / tuning aids, results are byte-identical whatever they
 * say):  RT_GROUP_K    bounce rays as grouped packets: (group radius / scene extent)^2, default 0.01
 *        RT_GROUP_COS  ... and the cosine of the largest angle between a group member and its leader, default 0.94
 *        RT_GROUPS_MIN object count from which the groups of 32 neighbouring objects are built, default 2048
 *        RT_BLOCK_LOG  path: the instrumented twin of the block kernel (rt_options.stats != NULL) dumps one 64-byte
 *                      timing record per block of 48 x 16 pixels there (tools/block_log.py) */
#define RT_HOST_STORE_AUTO   0
#define RT_HOST_STORE_ALWAYS 1
#define RT_HOST_STORE_NEVER  2

typedef struct rt_context rt_context;

/* Library/ABI introspection. */
int rt_abi_version(void);
const char* rt_last_error_string(void);
/* Number of kernels this library has launched in this process (bench evidence). */
uint64_t rt_kernel_launch_count(void);

/* Context: owns device buffers, a stream and pinned staging on one device.
 * Replaces the GL context + texture object (main.cpp:47-76, 152-159). */
int rt_create(rt_context** out_ctx, int device);
int rt_destroy(rt_context* ctx);
int rt_synchronize(rt_context* ctx);

void rt_options_default(rt_options* opts);

/* Page-locked host memory for out_rgba: lets the device->host copy of the frame
 * run at full PCIe rate (a pageable buffer works too, only slower). */
int rt_host_alloc(void** ptr, size_t bytes);
int rt_host_free(void* ptr);

/* THE drop-in call (north_star): render(scene, camera_uniforms, out_rgba, w, h).
 * Replaces glUseProgram+glBindImageTexture+glUniform1f+glDispatchCompute+glFinish
 * (main.cpp:220-238).  out_rgba: caller-owned host buffer of width*height*4
 * bytes, R,G,B,A tightly packed, row 0 = bottom scanline (GL order), A = 0
 * (glsl:404).  Synchronous like glFinish().  Uses a lazily created default
 * context on the current device. */
int rt_render(const rt_scene* scene, const rt_camera* camera_uniforms,
              uint8_t* out_rgba, int width, int height);

/* Extended call: explicit context and options (spp, row band, device output). */
int rt_render_ex(rt_context* ctx, const rt_scene* scene, const rt_camera* camera_uniforms,
                 void* out_rgba, int width, int height, const rt_options* opts);

/* Throughput mode (BASELINE config 5): n_frames cameras over one scene in ONE
 * launch; frame k is written at out_rgba + k*width*height*4. */
int rt_render_batch(rt_context* ctx, const rt_scene* scene, const rt_camera* cameras,
                    int n_frames, void* out_rgba, int width, int height, const rt_options* opts);

/* Two-phase use for steady-state rendering of a fixed scene: upload once
 * (host prologue: per-object matrices, glsl:650-652, 718), render many. */
int rt_upload_scene(rt_context* ctx, const rt_scene* scene);
int rt_render_uploaded(rt_context* ctx, const rt_camera* cameras, int n_frames,
                       void* out_rgba, int width, int height, const rt_options* opts);

/* rt_render_block_noalloc -- NO COUNTERPART IN THE REFERENCE.
 * For callers that must not allocate while rendering (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc): renders n_frames cameras over the scene already uploaded with
 * rt_upload_scene() into the caller's DEVICE buffer out_rgba_device (layout as rt_render_uploaded with
 * out_is_device = 1) through the CTA-level packet kernel (render_block_kernel), with all of that kernel's
 * scratch -- tile counter, camera records, per-CTA shadow-candidate lists -- in ONE file-scope device array of
 * RT_NOALLOC_SCRATCH_BYTES (8 MiB + 4 KiB; every GPU that loads the library holds it until process exit).
 * The call makes no cudaMalloc / cudaFree / cudaHostAlloc and calls no allocating library routine; the set-up
 * (rt_create, rt_upload_scene, the caller's output buffer) is done beforehand and may allocate.
 * Limits, rejected with RT_ERR_UNSUPPORTED and a message: RT_PACKET objects minimum (96 boxes + spheres: below
 * that the block kernel is not used at all), at most RT_NOALLOC_MAX_OBJECTS boxes + spheres, at most
 * RT_NOALLOC_MAX_FRAMES cameras; device output only; fast kernel only (opts->strict, opts->stats and the
 * no_cull switches 1 / 4 / 8 are refused).  Honoured options: spp, seed, y0 / y1, out_band_only, stream, async,
 * done_flag / done_value, no_cull bits 5-6.
 * The array is shared by every call on that GPU and keeps the last call's contents; what must start at zero
 * (the tile counter) is cleared by each call.  Calls that OVERLAP on the GPU -- two streams, two contexts on one
 * device, a second call before the first has finished under async -- race on it: serialise them.
 * Output: byte-identical to rt_render_uploaded on the same scene, cameras and options. */
#define RT_NOALLOC_MAX_OBJECTS 2048
#define RT_NOALLOC_MAX_FRAMES 16
#define RT_NOALLOC_SCRATCH_BYTES (4096 + 4 * 256 * 2048 * 4)
int rt_render_block_noalloc(rt_context* ctx, const rt_camera* cameras, int n_frames, void* out_rgba_device,
                            int width, int height, const rt_options* opts);

/* Where the per-object prologue of rt_upload_scene runs (the matrices the shader rebuilds per test,
 * glsl:650-652, 718): on the host, or as device kernels from the raw objects (large animated scenes:
 * 8.6 MB of objects over PCIe instead of 27 MB of records at 64k objects, no host trigonometry).
 * AUTO = device from RT_PROLOGUE_DEVICE_MIN_OBJECTS objects up.  The strict kernel always uses the
 * host run (a strict render after a device prologue re-runs it from a kept copy of the scene). */
#define RT_PROLOGUE_AUTO   0
#define RT_PROLOGUE_HOST   1
#define RT_PROLOGUE_DEVICE 2
#define RT_PROLOGUE_DEVICE_MIN_OBJECTS 512
int rt_set_prologue(rt_context* ctx, int mode);

/* The shader's hard-coded data as a function of its one real uniform `time`
 * (glsl:4): objects (glsl:261-321), lights (glsl:199-224), camera (glsl:334-364).
 * objects must have room for 5 rt_object, lights for 3 rt_light.  scene->max_depth
 * is set to the shipped MAX_RAYTRACE_DEPTH = 0 (glsl:22). */
int rt_builtin_scene(float time, rt_object* objects5, rt_light* lights3, rt_scene* scene);
int rt_builtin_camera(float time, rt_camera* camera);
/* Orbit camera from the shader's `speed` variable directly (glsl:343-358); used
 * by the turntable batch (config 5). */
int rt_orbit_camera(float speed, rt_camera* camera);

/* Scaled synthetic scene (BASELINE config 4; generator defined in SURVEY §8d):
 * object 0 is the builtin enclosing cube, objects 1..n-1 are oriented boxes
 * (spheres when i%5==4) placed by splitmix64(seed).  objects must hold n
 * rt_object, lights 3 rt_light. */
int rt_scaled_scene(int n_objects, uint64_t seed, rt_object* objects, rt_light* lights3, rt_scene* scene);

/* Scene / camera description on the wire: the data the shader hard-codes in its globals
 * (glsl:22, 74-157, 199-224, 261-321, 334-364) as a line-oriented text file, fields in the order
 * of the structs above, 9 significant digits (every FP32 bit pattern survives save -> load).
 * Format in openglraytracer_b200/csrc/rt_scene_io.cpp.  camera may be NULL on save; a file
 * without a camera record loads with rt_scene_file_camera() == NULL. */
typedef struct rt_scene_file rt_scene_file;
int rt_scene_save(const char* path, const rt_scene* scene, const rt_camera* camera);
int rt_scene_load(const char* path, rt_scene_file** out);
const rt_scene* rt_scene_file_scene(const rt_scene_file* sf);
const rt_camera* rt_scene_file_camera(const rt_scene_file* sf);
void rt_scene_file_free(rt_scene_file* sf);

/* Replaces the draw_screen_vert/frag blit (main.cpp:240-260): binary PPM (P6),
 * rows flipped so the picture is upright, alpha dropped (frag:9). */
int rt_write_ppm(const char* path, const uint8_t* rgba, int width, int height);
int rt_write_raw(const char* path, const uint8_t* rgba, int width, int height);

/* Microbenchmark: dependent-free FP32 FMA chains on every SM; returns measured
 * TFLOP/s (2 flop per FMA).  Denominator check for the FP32 roofline. */
int rt_fp32_peak_tflops(int device, int iters, double* tflops, float* ms);

/* Plain device memory on the current device (cudaMalloc / cudaFree): a framebuffer whose
 * CUDA IPC handle maps exactly onto the pointer (a pointer carved out of a caching allocator's
 * block would export the block). */
int rt_device_alloc(void** device_ptr, size_t bytes);
int rt_device_free(void* device_ptr);

/* Multi-GPU helper for the fused gather: export/import a CUDA IPC handle of a
 * device framebuffer so that other ranks' kernels store their row band directly
 * into rank 0's buffer over NVLink. handle is 64 bytes. */
int rt_ipc_export(void* device_ptr, uint8_t handle64[64]);
int rt_ipc_import(const uint8_t handle64[64], void** device_ptr);
int rt_ipc_close(void* device_ptr);

/* Consumer side of rt_options.done_flag (multi-GPU frame mode without a collective): one warp on `stream` waits
 * until each of the n words at flags_dev (device memory of the current device) has reached `value` (epochs only
 * grow; wrap-safe).  Work queued on the stream behind it sees every byte the producers stored before raising
 * their word.  Gives up after timeout_ms (<= 0: 2000) so that a lost producer cannot hang the GPU;
 * rt_wait_flags_status() then returns 1 (once: reading clears it). */
int rt_wait_flags(const uint32_t* flags_dev, int n, uint32_t value, int timeout_ms, void* stream);
int rt_wait_flags_status(void);
/* Raises a 32-bit word to `value` in stream order (system-scope release) -- the word may be read by other GPUs
 * that mapped the allocation ("frame consumed": producers wait for it before they overwrite the frame). */
int rt_write_flag(uint32_t* flag_dev, uint32_t value, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RT_B200_H */
