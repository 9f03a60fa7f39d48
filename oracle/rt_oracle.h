/*
 * rt_oracle.h — API of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  The oracle is a line-by-line FP32 restatement of
 * /root/reference/OpenGLRaytracer/raytrace_compute.glsl (glsl:1-1106) used as the
 * parity checker and as the timed CPU baseline.  Only tests/, bench.py's
 * cpu_baseline / --impl reference legs and __graft_entry__.smoke() may load it.
 * The product (openglraytracer_b200/, librt_b200.so) never links or calls it.
 *
 * PARITY UNPINNED: the reference has no tests, golden images or CPU
 * implementation of this path, and no GL/GLSL stack exists on this box to run
 * the shader itself, so there is nothing external to pin the restatement
 * against.  It is validated by review against glsl:N citations, analytic
 * known-answer tests and float64 sanity anchors (tests/test_oracle.py).
 *
 * Struct layouts come from include/rt_b200.h (the interface, not product code).
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H

#include "../include/rt_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_stats {
    uint64_t rays;          /* get_closest_collision calls (glsl:738) */
    uint64_t nodes;         /* stack frames processed (glsl:886) */
    uint64_t hit_nodes;
    uint64_t shadow_rays;   /* glsl:813 */
    uint64_t box_tests;     /* intersect_box_object calls (glsl:752) */
    uint64_t sphere_tests;  /* intersect_sphere_object calls (glsl:762) */
    uint64_t red_pixels;    /* runaway-guard samples (glsl:1098) */
    uint64_t loop_iterations; /* process_stack_element calls (glsl:1093) */
} oracle_stats;

typedef struct oracle_options {
    int32_t spp;            /* 1 = the shader as written */
    uint32_t seed;
    int32_t y0, y1;         /* row band; y1 <= 0 -> height */
    int32_t as_written;     /* 1 = redo per-pixel camera matrices and per-test object
                               matrices exactly like the shader (CPU-baseline mode);
                               0 = cache those pure functions once (bit-identical) */
    int32_t threads;        /* OpenMP threads; <= 0 -> all */
    const int32_t* pixel_list; /* optional (x,y) pairs; output is then n_pixels*4 bytes in list order */
    int32_t n_pixels;
    float* out_float;       /* optional: pre-quantisation RGB, 3 floats per output pixel */
    oracle_stats* stats;    /* optional */
    int32_t* out_nodes;     /* optional: bounce-tree nodes per output pixel (workload analysis) */
} oracle_options;

void oracle_options_default(oracle_options* o);

/* Full-frame (or band / pixel-list) render.  out_rgba layout as rt_render. */
int oracle_render(const rt_scene* scene, const rt_camera* camera, uint8_t* out_rgba,
                  int width, int height, const oracle_options* opts);

/* The shader's hard-coded data (glsl:199-224, 261-321, 334-364). */
void oracle_builtin_scene(float time, rt_object* objects5, rt_light* lights3, rt_scene* scene);
void oracle_builtin_camera(float time, rt_camera* camera);
void oracle_orbit_camera(float speed, rt_camera* camera);

/* Probes for known-answer tests. */
/* primary ray of pixel (glsl:377-392): dir[3]; closest hit: obj index (-1 miss), t, p[3], n[3], inside */
void oracle_probe_pixel(const rt_scene* scene, const rt_camera* camera, int width, int height,
                        int px, int py, float* dir3, int* obj, float* t, float* p3, float* n3, int* inside);
/* get_closest_collision for an arbitrary ray */
void oracle_probe_ray(const rt_scene* scene, const float* start3, const float* dir3,
                      int* obj, float* t, float* p3, float* n3, int* inside);
/* matrix helpers, column-major 16 floats */
void oracle_mat4_inverse(const float* m16, float* out16);
void oracle_mat4_mul(const float* a16, const float* b16, float* out16);
void oracle_transform_matrix(const float* pos3, const float* angles3, float* out16);
void oracle_projection_matrix(const rt_camera* c, float* out16);
void oracle_view_matrix(const rt_camera* c, float* out16);
void oracle_reflect(const float* i3, const float* n3, float* out3);
void oracle_refract(const float* i3, const float* n3, float eta, float* out3);
uint8_t oracle_unorm8(float f);
/* jitter offsets for sample s of pixel (x,y) */
void oracle_jitter(int x, int y, int s, int spp, uint32_t seed, float* jx, float* jy);
int oracle_num_threads(void);

#ifdef __cplusplus
}
#endif
#endif
