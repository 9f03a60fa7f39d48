/*
 * rt_device.h — device-side data layout shared by the host prologue (rt_host.cpp)
 * and the kernels (rt_kernel.cuh).
 *
 * HBM layout (everything is tiny next to 180 GB; what matters is the split):
 *   HOT  — what the brute-force loop (glsl:744-779) touches for EVERY (ray,
 *          object) pair.  Stored as dense float4 arrays, boxes and spheres in
 *          separate lists (each in original object order), so that a warp
 *          reads one record with 3 (fast) / 5 (strict) broadcast 16-byte loads
 *          from shared memory.
 *   COLD — what only the single closest hit of a ray needs (matrices for the
 *          hit point / normal, glsl:718-721, and the material, glsl:791).
 *          One 256-byte record per original object, read once per ray.
 */
#ifndef RT_DEVICE_H
#define RT_DEVICE_H

#include <stdint.h>
#include <vector_types.h>

/* same value as RT_MAX_DEPTH in include/rt_b200.h (checked in rt_host.cpp) */
#define RT_MAX_DEPTH_DEV 15

#define RT_KIND_SKIP   0
#define RT_KIND_BOX    1
#define RT_KIND_SPHERE 2

/* float4 per box in the hot list */
#define RT_BOX_F4_FAST   3   /* rows of the recentred, 1/half-extent-scaled world->local map */
#define RT_BOX_F4_STRICT 5   /* world->local rows (3) + (mins.xyz, maxs.x) + (maxs.y, maxs.z, 0, 0) */

/* from this many hot objects the large-scene kernels put the packet (cone) tests in front of the per-ray loops */
#define RT_PACKET_MIN_OBJECTS 96
/* from this many hot objects the CTA-level packet tests (rt_kernel_block.cuh) pay for their barriers */
#define RT_BLOCK_MIN_OBJECTS 128
/* with bounce rays (depth > 0) and this many hot objects the per-warp kernel is faster again: the bounce work of a
 * block's 8 work items is uneven and its warps wait for each other at the block's barriers (depth 2, round 2: 1k
 * objects 6.7 against 7.6 ms, 2k 9.6 / 9.7, 4k 17.2 / 15.0, 16k 50.8 / 39.3) */
#define RT_BLOCK_MAX_OBJECTS_BOUNCE 3072
/* groups of this many neighbouring objects carry a bounding sphere of their own (rt_groups.cu); built from
 * RT_GROUPS_MIN_OBJECTS hot objects up, for the fast kernel */
#define RT_GROUP_SIZE 32
#define RT_GROUPS_MIN_OBJECTS 2048
#define RT_GROUPS_MIN_OBJECTS_BOUNCE 4096      /* ... and used at depth > 0 only from this many */
#define RT_BLOCK_REC_CAP 8192      /* bounding records kept next to a CTA's shadow-candidate list (16 bytes each) */
#define RT_MAX_CTAS_PER_SM 4      /* upper bound of the persistent grid per SM (list slices of the block kernel) */
/* rt_render_block_noalloc: limits of the file-scope scratch array (same values as in include/rt_b200.h, checked in rt_host.cpp) */
#define RT_NOALLOC_MAX_OBJECTS_DEV 2048
#define RT_NOALLOC_MAX_FRAMES_DEV 16
#define RT_NOALLOC_SCRATCH_BYTES_DEV (4096 + 4 * 256 * 2048 * 4)
#ifndef RT_NOALLOC_MAX_OBJECTS
#define RT_NOALLOC_MAX_OBJECTS RT_NOALLOC_MAX_OBJECTS_DEV
#define RT_NOALLOC_MAX_FRAMES RT_NOALLOC_MAX_FRAMES_DEV
#define RT_NOALLOC_SCRATCH_BYTES RT_NOALLOC_SCRATCH_BYTES_DEV
#endif

/* Cold record: 64 words; the four material colours first so that each is one aligned 16-byte load. */
struct RtCold {
    /* material (glsl:56-69) */
    float ambient[4];
    float diffuse[4];
    float specular[4];
    float emissive[4];
    float w2l[12];      /* world->local rows: row i = (m[0][i], m[1][i], m[2][i], m[3][i]) of inverse(T*R), glsl:652 */
    float l2w[12];      /* local->world rows, glsl:650 */
    float nmat[9];      /* transpose(inverse(mat3(l2w))), column-major c[col][row], glsl:718 */
    float inv_ior;      /* 1 / refraction_index (glsl:1013), IEEE */
    float ior_back;     /* 1 / inv_ior (glsl:1017-1020: the ratio is inverted again for a ray leaving the object) */
    float unused_[4];
    int32_t kind;
    int32_t hot_index;  /* index in the box / sphere hot list */
    float shininess;
    float reflectivity;
    float transparency;
    float refraction_index;
    float position[3];  /* sphere centre */
};

struct RtLightDev {
    float position[4];  /* w = 1 if any diffuse/specular component is non-zero, else 0 */
    float ambient[4];
    float diffuse[4];
    float specular[4];
};

/* Per-frame camera uniforms after the host prologue: inverse(proj*view)
 * (glsl:383, pixel-invariant) and the ray origin (glsl:391). */
struct RtCameraDev {
    float inv_pv[16];   /* column-major c[col][row] */
    float position[4];
};

struct RtCounters {
    unsigned long long rays, nodes, hit_nodes, shadow_rays, box_tests, sphere_tests, red_pixels, pad;   /* pad = shadow rays resolved without traversal */
    unsigned long long cull_tests;   /* bounding-sphere tests executed for live rays (large-scene kernels) */
    unsigned long long warp_iters;   /* bounce-loop iterations summed over warps: nodes / (32 * warp_iters) = lane utilisation */
    unsigned long long packet_tests; /* (object, cone) tests of the packet loops, one per lane (large-scene kernels) */
    unsigned long long group_passes; /* per warp (lane 0): passes over the scene for a group of bounce rays / their hit points */
    unsigned long long loop_passes;  /* per warp (lane 0): per-ray loop passes for bounce rays the groups left over */
    unsigned long long candidates;   /* per warp (lane 0): candidate_test calls */
    unsigned long long shadow_group_passes;   /* per warp (lane 0): passes for the shadow rays of a group of hit points */
    unsigned long long shaded_lights;  /* (hit node, light) pairs whose diffuse + specular terms were evaluated (glsl:822-833) */
};

struct RtParams {
    const float4* box_hot;
    const float4* box_cull;     /* per box: (world-space centre, padded squared circumradius) */
    const float4* sph_hot;      /* (cx, cy, cz, r*r) */
    const float4* obj_lrec;     /* per 64-object tile, per lit light: (L - centre, |L - centre|^2 - R'^2), boxes then spheres */
    const int32_t* box_obj;     /* hot index -> original object index */
    const int32_t* sph_obj;
    const RtCold* cold;
    const RtLightDev* lights;
    const RtCameraDev* cams;
    uint8_t* out;               /* frame k at out + k*frame_stride */
    RtCounters* counters;       /* used by the instrumented twin only */
    unsigned long long* tile_counter;  /* dynamic tile scheduler, zeroed before each launch */
    long long frame_stride;     /* bytes */
    int32_t n_box, n_sph, n_lights, max_depth;
    int32_t n_frames, width, height, y0, y1;
    int32_t out_row0;           /* image row stored at out row 0 (y0 for band-only output, else 0) */
    int32_t spp, spp_n;         /* spp_n = sqrt(spp) if spp is a perfect square, else 0 */
    uint32_t seed;
    int32_t tiles_x, tiles_y;   /* 8x4-pixel warp tiles covering width x (y1-y0) */
    int32_t resident;           /* 1: whole hot scene staged in shared memory once per CTA */
    int32_t n_lit;              /* lit lights with per-light records (1..3), 0 = shadow rays take the generic path */
    int32_t n_lit_batch;        /* fast kernel: number of lit lights if <= 3 (one shadow batch, lit[] valid), else -1 */
    int32_t lit[3];             /* indices of the lit lights */
    float amb_sum[4];           /* sum of the lights' ambient terms, in light order (glsl:801) */
    int32_t cull;               /* 1: large-scene kernels run the slab test behind the bounding-sphere test */
    int32_t packet;             /* 1: coherent packets (primary rays, shadow rays per light) run the cone test first */
    int32_t block;              /* 1: CTA-level packet tests (blocks of 2 x 4 work items share the pass over the scene) */
    uint32_t* blist;            /* block kernel: per-CTA shadow-candidate lists, blist_stride entries each */
    int32_t blist_stride;
    float4* brec;               /* ... and, per CTA, the bounding records of the first brec_cap entries of its list (0: none) */
    int32_t brec_cap;
    const float* extent;        /* scene extent E (max |centre| + radius over objects and lights), written by the prologue */
    int32_t fan;                /* 1: shadow packets of the block kernel also carry a fan plane (rt_kernel.cuh: Cone::nrm) */
    float group_cos;            /* bounce rays join a group when their direction is within this cosine of the leader's */
    float group_k;              /* > 0: secondary rays are traced as grouped packets; group radius^2 = E^2 * group_k */
    uint32_t* blk_log;          /* block kernel, instrumented twin only (RT_BLOCK_LOG=path): 16 words per block, see tools/block_log.py */
    int32_t room_skip;          /* 1: hot box 0 holds every lit light (with margin): shadow rays that start inside it too skip it */
    float4 lit_pos[3], lit_dif[3], lit_spc[3];   /* the lit lights (n_lit_batch of them): position, diffuse, specular (glsl:190-196) */
    const float4* grp;          /* per group: (centre, radius) of a sphere around its members' bounding spheres (rt_groups.cu) */
    const float4* srec;         /* the bounding records (box_cull / sph_hot entries) in group order, 32 per group; w = NaN: none */
    const uint32_t* sidx;       /* their hot-list positions */
    int32_t n_groups;           /* 0: no groups (small scene, strict kernel, or switched off: no_cull bit 8) */
    int32_t split_cones;        /* 1: block kernel: a warp whose hit points spread wide as seen from a light gives them a cone per cluster */
    RtCameraDev cam0;           /* single-frame launches with cams == NULL: the camera uniforms ride in the kernel parameters */
};

#endif
