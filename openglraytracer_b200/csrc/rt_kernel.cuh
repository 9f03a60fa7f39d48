/*
 * rt_kernel.cuh — the per-pixel ray-scene kernel for sm_100a.
 *
 * Replaces raytrace_compute.glsl main() (glsl:325-405) and everything it calls.
 * Included twice: rt_kernel_fast.cu (RT_STRICT=0, default -fmad=true) and
 * rt_kernel_strict.cu (RT_STRICT=1, compiled with -fmad=false).
 *
 *   strict: the shader's arithmetic op by op (IEEE divide/sqrt, no FMA), used to
 *           prove logic equivalence with the oracle.
 *   fast:   the shipped kernel.  The ray/box slab test (glsl:655-666) runs in a
 *           recentred local frame scaled by 1/half-extent, so the six divides
 *           become 3 MUFU.RCP + 6 FFMA (t = -s*inv -+ |inv|); explicit
 *           __fmaf_rn/__fmul_rn keep the hot-loop t bit-reproducible when the
 *           winner is re-evaluated for its face index (glsl:698-704).
 *           No --use_fast_math; rcp/sqrt approximations are explicit and local
 *           to the hot loop (<= 2 ulp in t).
 *
 * Execution model (B200-first, not the shader's 1 thread per work-group):
 *   - persistent CTAs of 8 warps, 2 per SM; a warp owns 8x4-pixel tiles so
 *     primary rays stay coherent; the object loop is warp-uniform (brute force
 *     like glsl:744, no BVH), so there is no divergence inside it.
 *   - hot scene records staged once per CTA in shared memory and read by
 *     broadcast 16-byte LDS; cold records (hit matrices, material) read from
 *     global/L2 only for the closest hit.
 *   - the shader's 5-phase, 100-frame Stack_Element machine (glsl:848-1105) is
 *     replaced by a per-lane stack of pending rays carrying a scalar weight:
 *     mix() (glsl:1046,1050) is linear, so final = sum_nodes w(node)*phong(node).
 *   - shadow rays (glsl:807-819) are any-hit with t_max = 1 and are traced 3 at
 *     a time per lane (ILP + LDS amortisation), warp-vote early exit.
 *   - RGBA8 written with 16-byte vector stores (4 pixels gathered by shuffle).
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "rt_device.h"

#ifndef RT_STRICT
#define RT_STRICT 0
#endif

#if RT_STRICT
#define RT_NS rt_strict
#define RT_BOX_F4 RT_BOX_F4_STRICT
#else
#define RT_NS rt_fast
#define RT_BOX_F4 RT_BOX_F4_FAST
#endif

namespace RT_NS {

#ifndef RT_MIN_BLOCKS
#define RT_MIN_BLOCKS 2          // large scenes: the object loops are issue-bound, 16 warps/SM suffice
#endif
#ifndef RT_MIN_BLOCKS_CULL
#define RT_MIN_BLOCKS_CULL 2     // large scenes behind the bounding-sphere tests: short dependent chains on 16-byte LDS
#endif                           // results, latency-bound at 16 warps/SM -> 24 warps/SM (85 registers; shading spills a little)
#ifndef RT_MIN_BLOCKS_SMALL
#define RT_MIN_BLOCKS_SMALL 3    // small scenes: short loops, latency-bound shading -> 24 warps/SM
#endif
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr unsigned kFull = 0xffffffffu;
constexpr int kStack = RT_MAX_DEPTH_DEV + 1;
constexpr int kShadowBatch = 3;
constexpr float kFar = 10000.0f;     // glsl:740

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
// Everything that decides which object a ray hits or whether a shadow ray is blocked -- hit point, normal,
// shadow / bounce ray origins and directions -- is written with madd() so that every instantiation of the
// kernels (resident / streamed, with / without bounding tests, block kernel) rounds alike: left to the compiler,
// FMA contraction is chosen per instantiation, and one ulp in a hit point flips hit decisions in scenes where the
// shader's own quadratic is rounding noise (the +-3000 room of tests/test_gpu_parity.py).  Fast kernel: one
// FMA.  Strict kernel (-fmad=false): multiply then add, the shader's operations.
#if RT_STRICT
__device__ __forceinline__ float madd(float a, float b, float c) { return a * b + c; }
#else
__device__ __forceinline__ float madd(float a, float b, float c) { return __fmaf_rn(a, b, c); }
#endif
// Same association as the oracle's pinned dot(): ((x*x + y*y) + z*z)
__device__ __forceinline__ float dot3(V3 a, V3 b) { return madd(a.z, b.z, madd(a.y, b.y, __fmul_rn(a.x, b.x))); }
// a * s + b
__device__ __forceinline__ V3 axpy3(V3 a, float s, V3 b) { return mk(madd(a.x, s, b.x), madd(a.y, s, b.y), madd(a.z, s, b.z)); }
#if RT_STRICT
// GLSL normalize / divide / pow exactly as the oracle pins them (IEEE sqrt and divide, libm-grade powf)
__device__ __forceinline__ V3 normalize3(V3 v) { float l = sqrtf(dot3(v, v)); return mk(v.x / l, v.y / l, v.z / l); }
__device__ __forceinline__ float fdiv(float a, float b) { return a / b; }
__device__ __forceinline__ float fpow(float x, float y) { return powf(x, y); }
#else
// Shading-side arithmetic of the shipped kernel: v * rsqrt(v.v) (2 ulp), a * rcp(b) (2 ulp) and
// exp2(y * log2(x)) with the accurate log2f / exp2f (not the fast-math intrinsics).  These sit
// outside the hot loop but were ~40 % of the small-scene kernel's code size (instruction cache).
__device__ __forceinline__ V3 normalize3(V3 v) { float r = rsqrtf(dot3(v, v)); return mk(v.x * r, v.y * r, v.z * r); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdividef(a, b); }
// pow (glsl:832) as ex2.approx(y * lg2.approx(x)): two MUFU ops.  |error of lg2.approx| <= 2^-21.4 absolute on
// [0.5, 2), so the relative error of the result is about y * 4e-7 (shininess 10: 3e-6, against half an
// RGBA8 LSB = 2e-3); x = 0 gives ex2(-inf) = 0 and y = 0 gives 1 like powf.
__device__ __forceinline__ float fpow(float x, float y) {
    float l, r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(x));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(y * l));
    return (y == 0.0f) ? 1.0f : r;
}
#endif
__device__ __forceinline__ V3 reflect3(V3 I, V3 N) { float d2 = __fmul_rn(2.0f, dot3(N, I)); return axpy3(N, -d2, I); }   // I - d2 * N
__device__ __forceinline__ V3 refract3(V3 I, V3 N, float eta) {
    float ni = dot3(N, I);
    float k = madd(-__fmul_rn(eta, eta), madd(-ni, ni, 1.0f), 1.0f);          // 1 - eta * eta * (1 - ni * ni)
    if (k < 0.0f) return mk(0.0f, 0.0f, 0.0f);
    const float c = madd(eta, ni, sqrtf(k));
    return axpy3(N, -c, mk(__fmul_rn(eta, I.x), __fmul_rn(eta, I.y), __fmul_rn(eta, I.z)));      // eta * I - c * N
}
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float sqrt_approx(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

// -------------------------------------------------------------------------
// Ray / box slab test (glsl:655-666).  Returns t_near, t_far and (optionally)
// the per-axis boundaries for the face-index decision.
// -------------------------------------------------------------------------
struct Slabs { float t1x, t1y, t1z, t2x, t2y, t2z; };

#if RT_STRICT
// hot record: r0,r1,r2 = world->local rows, r3 = (mins.xyz, maxs.x), r4 = (maxs.y, maxs.z, -, -)
// (world_to_local * vec4(start,1)).xyz, glsl:655
__device__ __forceinline__ V3 box_origin(const float4* __restrict__ rec, V3 o) {
    const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2];
    V3 s;
    s.x = ((r0.x * o.x + r0.y * o.y) + r0.z * o.z) + r0.w * 1.0f;
    s.y = ((r1.x * o.x + r1.y * o.y) + r1.z * o.z) + r1.w * 1.0f;
    s.z = ((r2.x * o.x + r2.y * o.y) + r2.z * o.z) + r2.w * 1.0f;
    return s;
}
__device__ __forceinline__ void box_slabs(const float4* __restrict__ rec, V3 so, V3 d,
                                          float& tn, float& tf, Slabs& s, V3& ld) {
    const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2], r3 = rec[3], r4 = rec[4];
    // (world_to_local * vec4(dir,0)).xyz, glsl:656
    ld.x = ((r0.x * d.x + r0.y * d.y) + r0.z * d.z) + r0.w * 0.0f;
    ld.y = ((r1.x * d.x + r1.y * d.y) + r1.z * d.z) + r1.w * 0.0f;
    ld.z = ((r2.x * d.x + r2.y * d.y) + r2.z * d.z) + r2.w * 0.0f;
    float tminx = (r3.x - so.x) / ld.x, tmaxx = (r3.w - so.x) / ld.x;   // glsl:661-662
    float tminy = (r3.y - so.y) / ld.y, tmaxy = (r4.x - so.y) / ld.y;
    float tminz = (r3.z - so.z) / ld.z, tmaxz = (r4.y - so.z) / ld.z;
    s.t1x = fminf(tminx, tmaxx); s.t2x = fmaxf(tminx, tmaxx);           // glsl:663-664
    s.t1y = fminf(tminy, tmaxy); s.t2y = fmaxf(tminy, tmaxy);
    s.t1z = fminf(tminz, tmaxz); s.t2z = fmaxf(tminz, tmaxz);
    tn = fmaxf(fmaxf(s.t1x, s.t1y), s.t1z);                              // glsl:665
    tf = fminf(fminf(s.t2x, s.t2y), s.t2z);                              // glsl:666
}
// glsl:680: "the ray didn't intersect the box"
__device__ __forceinline__ bool box_reject(float tn, float tf) { return tn >= tf || tf <= 0.0f; }
#else
// hot record: rows of diag(1/h) * (world_to_local - centre): local'' = r * (p, 1), the box is [-1,1]^3
__device__ __forceinline__ V3 box_origin(const float4* __restrict__ rec, V3 o) {
    const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2];
    V3 s;
    s.x = __fmaf_rn(r0.x, o.x, __fmaf_rn(r0.y, o.y, __fmaf_rn(r0.z, o.z, r0.w)));
    s.y = __fmaf_rn(r1.x, o.x, __fmaf_rn(r1.y, o.y, __fmaf_rn(r1.z, o.z, r1.w)));
    s.z = __fmaf_rn(r2.x, o.x, __fmaf_rn(r2.y, o.y, __fmaf_rn(r2.z, o.z, r2.w)));
    return s;
}
// t = (+-1 - s)/d per axis as  n = -s*inv,  t1 = n - |inv|,  t2 = n + |inv|  (1 MUFU.RCP + 3 FP32 ops).
// The 1e-30 seed of the direction FMA chain keeps d'' away from an exact 0 (whose reciprocal would
// give inf - inf = NaN and lose the "parallel and outside the slab -> miss" decision, glsl:661-666)
// at no instruction cost; it is 2^-76 relative to any direction component that matters.
__device__ __forceinline__ void box_slabs(const float4* __restrict__ rec, V3 so, V3 d,
                                          float& tn, float& tf, Slabs& s, V3& ld) {
    const float4 r0 = rec[0], r1 = rec[1], r2 = rec[2];
    ld.x = __fmaf_rn(r0.x, d.x, __fmaf_rn(r0.y, d.y, __fmaf_rn(r0.z, d.z, 1e-30f)));
    ld.y = __fmaf_rn(r1.x, d.x, __fmaf_rn(r1.y, d.y, __fmaf_rn(r1.z, d.z, 1e-30f)));
    ld.z = __fmaf_rn(r2.x, d.x, __fmaf_rn(r2.y, d.y, __fmaf_rn(r2.z, d.z, 1e-30f)));
    const float ix = rcp_approx(ld.x), iy = rcp_approx(ld.y), iz = rcp_approx(ld.z);
    const float nx = __fmul_rn(-so.x, ix), ny = __fmul_rn(-so.y, iy), nz = __fmul_rn(-so.z, iz);
    s.t1x = __fadd_rn(nx, -fabsf(ix)); s.t2x = __fadd_rn(nx, fabsf(ix));
    s.t1y = __fadd_rn(ny, -fabsf(iy)); s.t2y = __fadd_rn(ny, fabsf(iy));
    s.t1z = __fadd_rn(nz, -fabsf(iz)); s.t2z = __fadd_rn(nz, fabsf(iz));
    tn = fmaxf(fmaxf(s.t1x, s.t1y), s.t1z);
    tf = fminf(fminf(s.t2x, s.t2y), s.t2z);
}
// All values are finite here, so "t_near >= t_far" (glsl:680) is exactly !(tn < tf); the t_far <= 0
// half of the reject is covered by the caller's t > 0 (glsl:753) because t = t_far when t_near < 0.
__device__ __forceinline__ bool box_reject(float tn, float tf) { return !(tn < tf); }
#endif
// glsl:690-696: the ray started inside the box -> the hit is the exit point
__device__ __forceinline__ float box_pick(float tn, float tf) { return (tn < 0.0f) ? tf : tn; }

// -------------------------------------------------------------------------
// Ray / sphere test (glsl:585-625).  c = (centre, r*r).  Split like the box test:
// the origin-dependent part (oc, qc) is shared by rays with a common origin.
// Returns t (or <= 0 / NaN for "no usable hit") and whether the origin is inside.
// -------------------------------------------------------------------------
struct RayQ { float qa, iqa; };        // per-ray invariants of the quadratic (fast path)
struct SphO { V3 oc; float qc; };      // per (origin, sphere)

__device__ __forceinline__ RayQ ray_q(V3 d) {
    RayQ q;
    q.qa = __fmaf_rn(d.x, d.x, __fmaf_rn(d.y, d.y, __fmul_rn(d.z, d.z)));
#if RT_STRICT
    q.iqa = 1.0f / q.qa;
#else
    q.iqa = rcp_approx(q.qa);       // 1 ulp in the sphere's t, like the box test's reciprocals
#endif
    return q;
}

#if RT_STRICT
__device__ __forceinline__ SphO sphere_origin(float4 c, V3 o) {
    SphO s;
    s.oc = mk(o.x - c.x, o.y - c.y, o.z - c.z);          // r.start - o.position
    s.qc = dot3(s.oc, s.oc) - c.w;                        // glsl:588
    return s;
}
__device__ __forceinline__ float sphere_t(SphO so, V3 d, RayQ, bool& inside) {
    float qa = dot3(d, d);                                // glsl:586
    float qb = dot3(2.0f * d, so.oc);                     // glsl:587
    float qd = qb * qb - 4.0f * qa * so.qc;               // glsl:591
    inside = false;
    if (qd < 0.0f) return -1.0f;
    float sq = sqrtf(qd);
    float t1 = (-qb + sq) / (2.0f * qa);
    float t2 = (-qb - sq) / (2.0f * qa);
    float tn = fminf(t1, t2), tf = fmaxf(t1, t2);
    if (tf < 0.0f) return -1.0f;
    if (tn < 0.0f) { inside = true; return tf; }
    return tn;
}
#else
__device__ __forceinline__ SphO sphere_origin(float4 c, V3 o) {
    SphO s;
    s.oc = mk(o.x - c.x, o.y - c.y, o.z - c.z);
    s.qc = __fmaf_rn(s.oc.x, s.oc.x, __fmaf_rn(s.oc.y, s.oc.y, __fmaf_rn(s.oc.z, s.oc.z, -c.w)));
    return s;
}
__device__ __forceinline__ float sphere_t(SphO so, V3 d, RayQ q, bool& inside) {
    float hb = __fmaf_rn(d.x, so.oc.x, __fmaf_rn(d.y, so.oc.y, __fmul_rn(d.z, so.oc.z)));   // qb / 2
    float disc = __fmaf_rn(hb, hb, -__fmul_rn(q.qa, so.qc));                                // qd / 4
    float sq = sqrt_approx(fmaxf(disc, 0.0f));
    float tn = __fmul_rn(-hb - sq, q.iqa);
    float tf = __fmul_rn(sq - hb, q.iqa);
    inside = tn < 0.0f;
    float t = inside ? tf : tn;
    return (disc < 0.0f) ? -1.0f : t;      // t_far < 0 gives t < 0: rejected by the caller's t > 0
}
#endif

// -------------------------------------------------------------------------
// Per-warp context: where this warp's object records come from, and its
// shared-memory scratch.
//   resident : the whole hot scene sits in the CTA's shared memory (staged once)
//   streaming: the hot scene is larger than shared memory; each warp pulls
//              kTileBox-record tiles through its own double buffer with TMA
//              bulk copies (cp.async.bulk + mbarrier complete_tx), next tile in
//              flight while the current one is tested.  Warps stay independent:
//              no CTA barrier inside the object loop.
// -------------------------------------------------------------------------
// All dynamic shared memory, addressed by float4 offset so that every record load is a plain LDS.128
// (a generic pointer picked at run time would compile to generic LD).
extern __shared__ float4 smem_f4[];

constexpr int kTileBox = 64;                                   // boxes per tile
constexpr int kTileBytes = kTileBox * RT_BOX_F4 * 16;
constexpr int kTileSph = kTileBytes / 16;                      // spheres per tile
constexpr int kRepackMinObjects = 48;                          // below this, shadow rays are never re-packed
constexpr int kQueueRays = 64;                                 // compacted shadow-ray queue entries
constexpr int kScratchBytes = kQueueRays * 8 * 4 + 96 * 4;     // queue + occlusion flags, per warp

struct WarpCtx {
    uint32_t box_off4, sph_off4; // resident: float4 offsets of the staged hot lists in smem_f4
    uint32_t lrec_off4;          // resident: float4 offset of the staged per-light records
    float* queue;                // [kQueueRays][8]
    int* occ;                    // [96]
    uint32_t buf_off4[2];        // streaming double buffer (float4 offsets in smem_f4)
    uint32_t buf_addr[2];        // shared-space addresses of buf / barriers
    uint32_t bar_addr[2];
    uint32_t phase;              // bit b: parity to wait for on barrier b
    int lane;
};

__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// 16-byte shared-memory load by 32-bit shared address.  volatile: the loads of one group of records
// stay where they are written, ahead of the arithmetic (ptxas otherwise issues each load right before
// its dependent chain and the loop runs at shared-memory latency).
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

// CULL: the box tiles hold 16-byte bounding-sphere records (kTileSph per tile); the 48-byte slab
// records stay in global memory / L2 and are read only when some lane passes the bounding-sphere test.
template <bool CULL> __device__ __forceinline__ constexpr int box_tile_n() { return CULL ? kTileSph : kTileBox; }
template <bool CULL> __device__ __forceinline__ constexpr int box_tile_f4() { return CULL ? 1 : RT_BOX_F4; }
template <bool CULL>
__device__ __forceinline__ int n_box_tiles(const RtParams& P) { return (P.n_box + box_tile_n<CULL>() - 1) / box_tile_n<CULL>(); }
__device__ __forceinline__ int n_sph_tiles(const RtParams& P) { return (P.n_sph + kTileSph - 1) / kTileSph; }

// lane 0: start the bulk copy of tile t (boxes first, then spheres) into buffer t & 1
template <bool CULL>
__device__ __forceinline__ void tile_issue(const RtParams& P, WarpCtx& w, int t) {
    const int nbt = n_box_tiles<CULL>(P);
    const void* src;
    uint32_t bytes;
    if (t < nbt) {
        const int first = t * box_tile_n<CULL>(), n = min(box_tile_n<CULL>(), P.n_box - first);
        src = CULL ? P.box_cull + first : P.box_hot + (size_t)first * RT_BOX_F4;
        bytes = (uint32_t)n * box_tile_f4<CULL>() * 16u;
    } else {
        const int first = (t - nbt) * kTileSph, n = min(kTileSph, P.n_sph - first);
        src = P.sph_hot + first;
        bytes = (uint32_t)n * 16u;
    }
    const int b = t & 1;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(w.bar_addr[b]), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(w.buf_addr[b]), "l"(src), "r"(bytes), "r"(w.bar_addr[b]) : "memory");
}
__device__ __forceinline__ void tile_wait(WarpCtx& w, int b) {
    const uint32_t parity = (w.phase >> b) & 1u;
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(w.bar_addr[b]), "r"(parity) : "memory");
    } while (!ok);
    w.phase ^= 1u << b;
}

// Returns the records of tile t.  Streaming: waits for its copy and puts tile t+1 in flight.
template <bool SMEM, bool CULL>
__device__ __forceinline__ const float4* tile_acquire(const RtParams& P, WarpCtx& w, int t, int n_tiles) {
    const int nbt = n_box_tiles<CULL>(P);
    if (SMEM) return smem_f4 + (t < nbt ? w.box_off4 + (uint32_t)t * box_tile_n<CULL>() * box_tile_f4<CULL>() : w.sph_off4 + (uint32_t)(t - nbt) * kTileSph);
    __syncwarp();                                     // every lane is done reading buffer (t+1)&1 (tile t-1)
    if (w.lane == 0 && t + 1 < n_tiles) tile_issue<CULL>(P, w, t + 1);
    tile_wait(w, t & 1);
    return smem_f4 + w.buf_off4[t & 1];
}
template <bool SMEM, bool CULL>
__device__ __forceinline__ void tile_begin(const RtParams& P, WarpCtx& w) {
    if (!SMEM && w.lane == 0) tile_issue<CULL>(P, w, 0);
}
// Early exit after tile t: the copy of tile t+1 is in flight; wait it out so the next trace starts clean.
template <bool SMEM>
__device__ __forceinline__ void tile_drain(WarpCtx& w, int t, int n_tiles) {
    if (!SMEM && t + 1 < n_tiles) tile_wait(w, (t + 1) & 1);
}

// -------------------------------------------------------------------------
// Brute-force loops over one tile (glsl:744-779).  Trip counts are warp-uniform.
// best[r] comes in as t_max (10000 closest-hit, 1 shadow, 0 inactive slot) and goes
// out as the smallest accepted t; bidx[r] = hot-list position of the winner
// (boxes 0..n_box-1, spheres n_box..).
//   SHARED : all R rays start at o[0] (the shadow rays of one hit point, glsl:808):
//            the origin is moved into the primitive's frame once per primitive.
//   ANYHIT : shadow semantics (glsl:816: in shadow iff ANY object is hit at 0 < t < 1).
// Most tests miss, so the accept logic sits behind one warp vote per box.
// -------------------------------------------------------------------------
template <int R, bool ANYHIT, bool SHARED, bool COUNT, bool VOTE = true>
__device__ __forceinline__ void boxes_tile(const float4* __restrict__ recs, int n, int base,
                                           const V3 (&o)[kShadowBatch], const V3 (&d)[kShadowBatch],
                                           float (&best)[kShadowBatch], int (&bidx)[kShadowBatch],
                                           unsigned long long& n_tests) {
#pragma unroll 2
    for (int i = 0; i < n; ++i) {
        float4 rec[RT_BOX_F4];
#pragma unroll
        for (int k = 0; k < RT_BOX_F4; ++k) rec[k] = recs[i * RT_BOX_F4 + k];
        float tn[R], tf[R];
        bool cand = false;
        V3 so;
        if (SHARED) so = box_origin(rec, o[0]);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (COUNT) n_tests += ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f);
            if (!SHARED) so = box_origin(rec, o[r]);
            Slabs s; V3 ld;
            box_slabs(rec, so, d[r], tn[r], tf[r], s, ld);
            cand = cand || !box_reject(tn[r], tf[r]);
        }
        // VOTE = false (a handful of large objects, e.g. the built-in scene: most tests hit): the accept
        // logic runs straight-line, the vote and branch would only add to it
        if (!VOTE || __any_sync(kFull, cand)) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float t = box_pick(tn[r], tf[r]);
                if (!box_reject(tn[r], tf[r]) && t > 0.0f && t < best[r]) { best[r] = t; if (!ANYHIT) bidx[r] = base + i; }
            }
        }
    }
}

// -------------------------------------------------------------------------
// Boxes behind a conservative bounding-sphere test (large scenes).
// recs[i] = (centre of box base+i in world space, padded squared circumradius).  A ray's LINE
// misses that sphere iff hb^2 < qa*qc (negative discriminant); the slab test of glsl:655-680 can
// only succeed for a line that meets the box, hence its circumsphere, so skipping it then changes
// nothing.  The radius is padded by 1e-3 (relative, host prologue) and qa by 1e-4, orders of
// magnitude more than the FP32 rounding of either test, so the decision is conservative and the
// image is byte-identical with the test on or off (tests/test_gpu_parity.py).  Still brute force:
// every ray meets every object (glsl:744), in original order.  Cost: 5 FP32 instructions per
// (ray, box) instead of 21; the 48-byte slab record is read from global/L2 only behind the vote.
// -------------------------------------------------------------------------
// boxes_tile for the large-scene kernels without the bounding-sphere test (rt_options.no_cull, and the
// strict kernel's large scenes): two boxes per iteration, the records of the next two software-
// pipelined (shared-memory latency overlaps the 40-odd instructions of the current pair).
template <int R, bool ANYHIT, bool SHARED, bool COUNT>
__device__ __forceinline__ void boxes_tile_pipe(const float4* __restrict__ recs, int n, int base,
                                                const V3 (&o)[kShadowBatch], const V3 (&d)[kShadowBatch],
                                                float (&best)[kShadowBatch], int (&bidx)[kShadowBatch],
                                                unsigned long long& n_tests) {
    constexpr int G = 2;
    constexpr uint32_t kRecBytes = RT_BOX_F4 * 16u;
    const uint32_t ra = smem_addr(recs);
    float4 nx[G][RT_BOX_F4];
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
        for (int k = 0; k < RT_BOX_F4; ++k) nx[g][k] = lds128(ra + (uint32_t)g * kRecBytes + (uint32_t)k * 16u);
    }
    for (int i = 0; i < n; i += G) {
        float4 rec[G][RT_BOX_F4];
        const uint32_t nxt = (uint32_t)(i + G < n ? i + G : i) * kRecBytes;     // last pair: a harmless re-load
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
            for (int k = 0; k < RT_BOX_F4; ++k) {
                rec[g][k] = nx[g][k];
                nx[g][k] = lds128(ra + nxt + (uint32_t)g * kRecBytes + (uint32_t)k * 16u);
            }
        }
        float tn[G][R], tf[G][R];
        bool cand = false;
#pragma unroll
        for (int g = 0; g < G; ++g) {
            V3 so;
            if (SHARED) so = box_origin(rec[g], o[0]);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (COUNT) n_tests += (i + g < n) && (ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f));
                if (!SHARED) so = box_origin(rec[g], o[r]);
                Slabs s; V3 ld;
                box_slabs(rec[g], so, d[r], tn[g][r], tf[g][r], s, ld);
                cand = cand || !box_reject(tn[g][r], tf[g][r]);
            }
        }
        if (__any_sync(kFull, cand)) {
#pragma unroll
            for (int g = 0; g < G; ++g) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float t = box_pick(tn[g][r], tf[g][r]);
                    // i + g < n: the pair may reach one record past the list (padding / the next list)
                    if (i + g < n && !box_reject(tn[g][r], tf[g][r]) && t > 0.0f && t < best[r]) {
                        best[r] = t;
                        if (!ANYHIT) bidx[r] = base + i + g;
                    }
                }
            }
        }
    }
}

// The slab test of box `idx` for the rays whose bit is set in pm (rare path, kept out of the unrolled loop).
template <int R, bool ANYHIT, bool SHARED, bool COUNT>
__device__ __forceinline__ void box_full_test(const RtParams& P, int idx, unsigned pm,
                                              const V3 (&o)[kShadowBatch], const V3 (&d)[kShadowBatch],
                                              float (&best)[kShadowBatch], int (&bidx)[kShadowBatch], RtCounters& cnt) {
    const float4* __restrict__ g = P.box_hot + (size_t)idx * RT_BOX_F4;
    float4 rec[RT_BOX_F4];
#pragma unroll
    for (int k = 0; k < RT_BOX_F4; ++k) rec[k] = __ldg(g + k);
    V3 bo;
    if (SHARED) bo = box_origin(rec, o[0]);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const bool mine = (pm >> r) & 1u;
        if (R > 1 && !__any_sync(kFull, mine)) continue;      // warp-uniform: no lane wants this ray slot tested
        if (COUNT) cnt.box_tests += mine;
        if (!SHARED) bo = box_origin(rec, o[r]);
        float tn, tf; Slabs s; V3 ld;
        box_slabs(rec, bo, d[r], tn, tf, s, ld);
        const float t = box_pick(tn, tf);
        // `mine` keeps the outcome a function of the ray alone (not of which lanes share its warp)
        if (mine && !box_reject(tn, tf) && t > 0.0f && t < best[r]) { best[r] = t; if (!ANYHIT) bidx[r] = idx; }
    }
}

template <int R, bool ANYHIT, bool SHARED, bool COUNT>
__device__ __forceinline__ void boxes_tile_cull(const RtParams& P, const float4* __restrict__ recs, int n, int base,
                                                const V3 (&o_in)[kShadowBatch], const V3 (&d_in)[kShadowBatch],
                                                float (&best_io)[kShadowBatch], int (&bidx_io)[kShadowBatch],
                                                RtCounters& cnt) {
    // Out of line on purpose (see lights_tile): own register allocation, loads issued per group.
    V3 o[kShadowBatch], d[kShadowBatch];
    float best[kShadowBatch]; int bidx[kShadowBatch];
#pragma unroll
    for (int r = 0; r < kShadowBatch; ++r) { o[r] = o_in[r]; d[r] = d_in[r]; best[r] = best_io[r]; bidx[r] = bidx_io[r]; }
    constexpr int G = R == 1 ? 8 : 4;    // boxes per warp vote; G record loads in flight
    float qe[R], qf[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        qe[r] = __fmul_rn(__fmaf_rn(d[r].x, d[r].x, __fmaf_rn(d[r].y, d[r].y, __fmul_rn(d[r].z, d[r].z))), 0.9999f);
        // Fast path: G boxes, one predicate, one vote.  A ray that is no longer live (occluded shadow
        // ray, idle lane) gets qa = 1e30 so that it fails the test unless its origin lies inside the
        // sphere; the slow path below re-checks liveness exactly, so that only costs a rare detour.
        qf[r] = (ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f)) ? qe[r] : 1e30f;   // refreshed after every slow path
    }
    const uint32_t ra = smem_addr(recs);
    // Software pipeline: the records of group i + G are loaded while group i is tested.  A record
    // past n (read from the padding / the next list) can only cause a detour through the slow path.
    float4 cn[G];
#pragma unroll
    for (int g = 0; g < G; ++g) cn[g] = lds128(ra + (uint32_t)g * 16u);
    for (int i = 0; i < n; i += G) {
        if (COUNT) {
#pragma unroll
            for (int r = 0; r < R; ++r)
                cnt.cull_tests += (ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f)) ? (unsigned long long)min(G, n - i) : 0ull;
        }
        float4 c[G];
        const uint32_t nxt = (uint32_t)(i + G < n ? i + G : i) * 16u;      // last group: a harmless re-load
#pragma unroll
        for (int g = 0; g < G; ++g) { c[g] = cn[g]; cn[g] = lds128(ra + nxt + (uint32_t)g * 16u); }
        bool any = false;
#pragma unroll
        for (int g = 0; g < G; ++g) {
            SphO so;
            if (SHARED) so = sphere_origin(c[g], o[0]);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (!SHARED) so = sphere_origin(c[g], o[r]);
                const float hb = __fmaf_rn(d[r].x, so.oc.x, __fmaf_rn(d[r].y, so.oc.y, __fmul_rn(d[r].z, so.oc.z)));
                const float disc = __fmaf_rn(hb, hb, -__fmul_rn(qf[r], so.qc));
                any = any || disc >= 0.0f;
            }
        }
        if (__any_sync(kFull, any)) {
            // Slow path (about one group in ten): redo the G tests one box at a time, exactly.
#pragma unroll 1
            for (int g = 0; g < G && i + g < n; ++g) {
                const float4 c = recs[i + g];
                unsigned m = 0u;
                SphO so;
                if (SHARED) so = sphere_origin(c, o[0]);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (!SHARED) so = sphere_origin(c, o[r]);
                    const bool live = ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f);
                    const float hb = __fmaf_rn(d[r].x, so.oc.x, __fmaf_rn(d[r].y, so.oc.y, __fmul_rn(d[r].z, so.oc.z)));
                    const float disc = __fmaf_rn(hb, hb, -__fmul_rn(qe[r], so.qc));
                    if (live && disc >= 0.0f) m |= 1u << r;
                }
                if (__any_sync(kFull, m != 0u))
                    box_full_test<R, ANYHIT, SHARED, COUNT>(P, base + i + g, m, o, d, best, bidx, cnt);
            }
#pragma unroll
            for (int r = 0; r < R; ++r) qf[r] = (ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f)) ? qe[r] : 1e30f;
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) { best_io[r] = best[r]; bidx_io[r] = bidx[r]; }
}

template <int R, bool ANYHIT, bool SHARED, bool COUNT, bool CULL>
__device__ __forceinline__ void spheres_tile(const RtParams& P, const float4* __restrict__ recs, int n, int base,
                                             const V3 (&o)[kShadowBatch], const V3 (&d)[kShadowBatch],
                                             float (&best)[kShadowBatch], int (&bidx)[kShadowBatch],
                                             RtCounters& cnt) {
    RayQ q[R];
#pragma unroll
    for (int r = 0; r < R; ++r) q[r] = ray_q(d[r]);
#if !RT_STRICT
    if (CULL) {
        // Large scenes: the discriminant of glsl:591 first (5 FP32 instructions per ray), G spheres per
        // warp vote, records software-pipelined like boxes_tile_cull; the roots are computed only for a
        // sphere some lane has a non-negative discriminant for.  Same expressions as sphere_t(),
        // evaluated in two steps, so the result is identical.  A dead ray tests with qa = 1e30.
        constexpr int G = R == 1 ? 8 : 4;
        float qf[R];
#pragma unroll
        for (int r = 0; r < R; ++r) qf[r] = (ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f)) ? q[r].qa : 1e30f;
        const uint32_t ra = smem_addr(recs);
        float4 cn[G];
#pragma unroll
        for (int g = 0; g < G; ++g) cn[g] = lds128(ra + (uint32_t)g * 16u);
        for (int j0 = 0; j0 < n; j0 += G) {
            if (COUNT) {
#pragma unroll
                for (int r = 0; r < R; ++r)
                    cnt.cull_tests += (ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f)) ? (unsigned long long)min(G, n - j0) : 0ull;
            }
            float4 c[G];
            const uint32_t nxt = (uint32_t)(j0 + G < n ? j0 + G : j0) * 16u;
#pragma unroll
            for (int g = 0; g < G; ++g) { c[g] = cn[g]; cn[g] = lds128(ra + nxt + (uint32_t)g * 16u); }
            bool any = false;
#pragma unroll
            for (int g = 0; g < G; ++g) {
                SphO so;
                if (SHARED) so = sphere_origin(c[g], o[0]);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (!SHARED) so = sphere_origin(c[g], o[r]);
                    const float hb = __fmaf_rn(d[r].x, so.oc.x, __fmaf_rn(d[r].y, so.oc.y, __fmul_rn(d[r].z, so.oc.z)));
                    const float disc = __fmaf_rn(hb, hb, -__fmul_rn(qf[r], so.qc));
                    any = any || disc >= 0.0f;
                }
            }
            if (__any_sync(kFull, any)) {
#pragma unroll 1
                for (int g = 0; g < G && j0 + g < n; ++g) {
                    const int j = j0 + g;
                    const float4 cc = recs[j];
                    SphO so;
                    if (SHARED) so = sphere_origin(cc, o[0]);
                    float hb[R], disc[R];
                    bool anyg = false;
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        if (!SHARED) so = sphere_origin(cc, o[r]);
                        const bool live = ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f);
                        hb[r] = __fmaf_rn(d[r].x, so.oc.x, __fmaf_rn(d[r].y, so.oc.y, __fmul_rn(d[r].z, so.oc.z)));
                        disc[r] = __fmaf_rn(hb[r], hb[r], -__fmul_rn(q[r].qa, so.qc));
                        anyg = anyg || (live && disc[r] >= 0.0f);
                    }
                    if (__any_sync(kFull, anyg)) {
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            if (COUNT) cnt.sphere_tests += (ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f)) && disc[r] >= 0.0f;
                            const float sq = sqrt_approx(fmaxf(disc[r], 0.0f));
                            const float tn = __fmul_rn(-hb[r] - sq, q[r].iqa);
                            const float tf = __fmul_rn(sq - hb[r], q[r].iqa);
                            const float t = (disc[r] < 0.0f) ? -1.0f : ((tn < 0.0f) ? tf : tn);
                            if (t > 0.0f && t < best[r]) { best[r] = t; if (!ANYHIT) bidx[r] = P.n_box + base + j; }
                            else if (!ANYHIT && t == best[r] && bidx[r] >= 0 && bidx[r] < P.n_box) {
                                if (__ldg(P.sph_obj + base + j) < __ldg(P.box_obj + bidx[r])) bidx[r] = P.n_box + base + j;
                            }
                        }
                    }
                }
#pragma unroll
                for (int r = 0; r < R; ++r) qf[r] = (ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f)) ? q[r].qa : 1e30f;
            }
        }
        return;
    }
#endif
#pragma unroll 2
    for (int j = 0; j < n; ++j) {
        const float4 c = recs[j];
        SphO so;
        if (SHARED) so = sphere_origin(c, o[0]);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (COUNT) cnt.sphere_tests += ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f);
            if (!SHARED) so = sphere_origin(c, o[r]);
            bool inside;
            const float t = sphere_t(so, d[r], q[r], inside);
            if (t > 0.0f && t < best[r]) { best[r] = t; if (!ANYHIT) bidx[r] = P.n_box + base + j; }
            else if (!ANYHIT && t == best[r] && bidx[r] >= 0 && bidx[r] < P.n_box) {
                // exact tie with a box: the lower ORIGINAL index wins (glsl:773 strict '<' in index order)
                if (__ldg(P.sph_obj + base + j) < __ldg(P.box_obj + bidx[r])) bidx[r] = P.n_box + base + j;
            }
        }
    }
}

// -------------------------------------------------------------------------
// Shadow rays against per-light records (large scenes, 1..3 lit lights).
// A shadow ray runs from p + 0.01 n along L - p (glsl:808-809): its line is the line
// through p and the light L, shifted by 0.01 n.  So "the line through L along d meets
// the primitive's bounding sphere grown by 0.0101" is a conservative test for "the ray's
// line meets the bounding sphere", and it needs no per-pixel origin work: the host stores,
// for every hot object (boxes then spheres) and every lit light, (L - c, |L - c|^2 - R'^2).
// Per (ray, object): one 16-byte LDS + 3 FFMA/FMUL (d . oc) + 2 (discriminant) + compare,
// against 21 instructions for the slab test; the exact test runs behind the warp vote.
// Tiles hold kTileBox objects x n_lit records: tile t = [light 0: 64][light 1: 64]...
// -------------------------------------------------------------------------
__device__ __forceinline__ int n_lrec_tiles(const RtParams& P) { return (P.n_box + P.n_sph + kTileBox - 1) / kTileBox; }

__device__ __forceinline__ void lrec_issue(const RtParams& P, WarpCtx& w, int t) {
    const uint32_t bytes = (uint32_t)P.n_lit * kTileBox * 16u;
    const void* src = P.obj_lrec + (size_t)t * P.n_lit * kTileBox;
    const int b = t & 1;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(w.bar_addr[b]), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(w.buf_addr[b]), "l"(src), "r"(bytes), "r"(w.bar_addr[b]) : "memory");
}
template <bool SMEM>
__device__ __forceinline__ const float4* lrec_acquire(const RtParams& P, WarpCtx& w, int t, int n_tiles) {
    if (SMEM) return smem_f4 + w.lrec_off4 + (uint32_t)t * P.n_lit * kTileBox;
    __syncwarp();
    if (w.lane == 0 && t + 1 < n_tiles) lrec_issue(P, w, t + 1);
    tile_wait(w, t & 1);
    return smem_f4 + w.buf_off4[t & 1];
}

// exact sphere test (glsl:585-625) of hot sphere j for the rays in m (rare path)
template <int R, bool COUNT>
__device__ __forceinline__ void sphere_full_test(const RtParams& P, int j, unsigned m,
                                                 const V3 (&o)[kShadowBatch], const V3 (&d)[kShadowBatch],
                                                 float (&best)[kShadowBatch], RtCounters& cnt) {
    const float4 c = __ldg(P.sph_hot + j);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const bool mine = (m >> r) & 1u;
        if (COUNT) cnt.sphere_tests += mine;
        bool inside;
        const float t = sphere_t(sphere_origin(c, o[r]), d[r], ray_q(d[r]), inside);
        if (mine && t > 0.0f && t < best[r]) best[r] = t;
    }
}

template <int R, bool COUNT>
__device__ __forceinline__ void lights_tile(const RtParams& P, const float4* __restrict__ recs, int t,
                                            const V3 (&o_in)[kShadowBatch], const V3 (&d_in)[kShadowBatch],
                                            float (&best_io)[kShadowBatch], const int (&src)[kShadowBatch],
                                            RtCounters& cnt) {
    // Out of line on purpose: the caller's shading state (some 60 live values) stays out of this
    // function's register allocation, so the record loads of a group are issued together.  The rays
    // are copied into registers here (the reference parameters live in the caller's stack frame).
    V3 o[kShadowBatch], d[kShadowBatch];
    float best[kShadowBatch];
#pragma unroll
    for (int r = 0; r < kShadowBatch; ++r) { o[r] = o_in[r]; d[r] = d_in[r]; best[r] = best_io[r]; }
    constexpr int G = R == 1 ? 8 : 4;          // objects per warp vote; G * R record loads in flight
    const int first = t * kTileBox;
    const int n = min(kTileBox, P.n_box + P.n_sph - first);
    float qe[R], qf[R];
    uint32_t sec[R];              // float4 offset of the ray's light section in the tile (slot = src % 3)
    uint32_t ra[R];               // shared address of that section
    const uint32_t base = smem_addr(recs);
#pragma unroll
    for (int r = 0; r < R; ++r) {
        qe[r] = __fmul_rn(__fmaf_rn(d[r].x, d[r].x, __fmaf_rn(d[r].y, d[r].y, __fmul_rn(d[r].z, d[r].z))), 0.9999f);
        const int slot = src[r] >= 0 ? src[r] % kShadowBatch : 0;
        sec[r] = (uint32_t)min(slot, P.n_lit - 1) * kTileBox;
        ra[r] = base + sec[r] * 16u;
        qf[r] = best[r] >= 1.0f ? qe[r] : 1e30f;     // refreshed after every slow path (the only place a ray dies)
    }
    // Software pipeline: the records of group i + G are loaded while group i is tested (the loads
    // cannot sink below the loop's back edge, so they overlap the arithmetic of the current group).
    float4 cn[G][R];                                 // (L - centre, |L - centre|^2 - R'^2); the tiles are padded
#pragma unroll
    for (int g = 0; g < G; ++g) {
#pragma unroll
        for (int r = 0; r < R; ++r) cn[g][r] = lds128(ra[r] + (uint32_t)g * 16u);
    }
    for (int i = 0; i < n; i += G) {
        if (COUNT) {
#pragma unroll
            for (int r = 0; r < R; ++r) cnt.cull_tests += best[r] >= 1.0f ? (unsigned long long)min(G, n - i) : 0ull;
        }
        float4 c[G][R];
        const uint32_t nxt = (uint32_t)min(i + G, kTileBox - G) * 16u;     // last group: a harmless re-load
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
            for (int r = 0; r < R; ++r) { c[g][r] = cn[g][r]; cn[g][r] = lds128(ra[r] + nxt + (uint32_t)g * 16u); }
        }
        bool any = false;
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float hb = __fmaf_rn(d[r].x, c[g][r].x, __fmaf_rn(d[r].y, c[g][r].y, __fmul_rn(d[r].z, c[g][r].z)));
                const float disc = __fmaf_rn(hb, hb, -__fmul_rn(qf[r], c[g][r].w));
                any = any || disc >= 0.0f;
            }
        }
        if (__any_sync(kFull, any)) {
#pragma unroll 1
            for (int g = 0; g < G && i + g < n; ++g) {
                unsigned m = 0u;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float4 c = recs[sec[r] + i + g];
                    const float hb = __fmaf_rn(d[r].x, c.x, __fmaf_rn(d[r].y, c.y, __fmul_rn(d[r].z, c.z)));
                    const float disc = __fmaf_rn(hb, hb, -__fmul_rn(qe[r], c.w));
                    if (best[r] >= 1.0f && disc >= 0.0f) m |= 1u << r;
                }
                if (__any_sync(kFull, m != 0u)) {
                    const int idx = first + i + g;
                    int unused[kShadowBatch];
                    if (idx < P.n_box) box_full_test<R, true, false, COUNT>(P, idx, m, o, d, best, unused, cnt);
                    else sphere_full_test<R, COUNT>(P, idx - P.n_box, m, o, d, best, cnt);
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) qf[r] = best[r] >= 1.0f ? qe[r] : 1e30f;
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) best_io[r] = best[r];
}

// -------------------------------------------------------------------------
// Packet tests (large scenes, coherent rays).  The rays of a warp that leave one point -- the primary
// rays of a work item (camera, glsl:391) or the shadow rays toward one light seen from the light
// (glsl:808-809) -- lie inside a narrow cone.  The warp tests 32 objects per instruction against that
// cone, one object per LANE (centre and padded bounding radius from the same 16-byte records), and
// only the objects whose bounding sphere meets the cone go through the per-ray tests above, in
// original order.  Still brute force -- every object of the scene is tested for every packet, no
// hierarchy, no spatial sorting -- but one cone test stands for 96 (ray, object) bounding tests.
//
// Conservative by construction: a ray of the packet can only hit an object
// whose bounding sphere, grown by `rp`'s padding, meets the cone: signed distance of the centre to the
// cone's surface, p*cos(theta) - h*sin(theta) <= rp, with sin(theta) padded by 5e-4 relative + 4e-6
// and p^2 = |v|^2 - h^2 lowered by its rounding bound.  The per-ray tests behind it are the ones of the loops
// above, so a frame is byte-identical with the packet tests on or off (tests/test_gpu_parity.py).
// -------------------------------------------------------------------------
#include "rt_cone.h"

__device__ __forceinline__ float warp_max_nonneg(float v) {      // v >= 0: floats order like their bit patterns
    return __uint_as_float(__reduce_max_sync(kFull, __float_as_uint(v)));
}
constexpr float kConeMinCos = 0.6f;     // wider packets take the per-ray loops

// The three rays of a lane going through the candidate tests.
struct RaySet {
    V3 o[kShadowBatch], d[kShadowBatch];
    float best[kShadowBatch];
    int bidx[kShadowBatch];
    float qe[kShadowBatch];           // |d|^2 * 0.9999 for the bounding test
    RayQ q[kShadowBatch];
};
__device__ __forceinline__ void rayset_load(RaySet& rs, const V3* __restrict__ o_in, const V3* __restrict__ d_in,
                                            const float* __restrict__ best_in, const int* __restrict__ bidx_in) {
#pragma unroll
    for (int r = 0; r < kShadowBatch; ++r) {
        rs.o[r] = o_in[r]; rs.d[r] = d_in[r]; rs.best[r] = best_in[r]; rs.bidx[r] = bidx_in[r];
        rs.q[r] = ray_q(rs.d[r]);
        rs.qe[r] = __fmul_rn(rs.q[r].qa, 0.9999f);
    }
}
// original object index of a hot-list position (boxes 0..n_box-1, then spheres)
__device__ __forceinline__ int hot_object(const RtParams& P, int hot) {
    return hot < P.n_box ? __ldg(P.box_obj + hot) : __ldg(P.sph_obj + (hot - P.n_box));
}
// One candidate object (hot-list position `hot`, bounding record cc) against the warp's rays: per-ray
// bounding-sphere test, then the exact test behind a warp vote (the slow paths of boxes_tile_cull /
// spheres_tile).  Accept rule: smaller t wins; on an exact tie the lower ORIGINAL object index wins.
// That is what visiting the objects in index order with a strict '<' gives (glsl:744-776), written so
// that the outcome does not depend on the order the candidates arrive in.
//   shared (warp-uniform): the three rays of a lane start at one point, one origin transform
//   want: ray slots this candidate is relevant for (bit r); others are skipped
template <bool COUNT>
__device__ __forceinline__ void candidate_test(const RtParams& P, RaySet& rs, int hot, float4 cc, bool shared, float live_min,
                                               unsigned want, RtCounters& cnt) {
    constexpr int R = kShadowBatch;
    const bool anyhit = live_min >= 1.0f;
    SphO so;
    if (COUNT && (threadIdx.x & 31) == 0) cnt.candidates++;
    if (hot < P.n_box) {
        unsigned m = 0u;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (r > 0 && !((want >> r) & 1u)) continue;       // warp-uniform (one ray per lane: slot 0 only)
            if (r == 0 || !shared) so = sphere_origin(cc, rs.o[r]);
            const bool live = rs.best[r] >= live_min && ((want >> r) & 1u);
            if (COUNT) cnt.cull_tests += live;
            const float hb = __fmaf_rn(rs.d[r].x, so.oc.x, __fmaf_rn(rs.d[r].y, so.oc.y, __fmul_rn(rs.d[r].z, so.oc.z)));
            const float disc = __fmaf_rn(hb, hb, -__fmul_rn(rs.qe[r], so.qc));
            if (live && disc >= 0.0f) m |= 1u << r;
        }
        if (!__any_sync(kFull, m != 0u)) return;
        const float4* __restrict__ g = P.box_hot + (size_t)hot * RT_BOX_F4;
        float4 rec[RT_BOX_F4];
#pragma unroll
        for (int k = 0; k < RT_BOX_F4; ++k) rec[k] = __ldg(g + k);
        V3 bo;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const bool mine = (m >> r) & 1u;
            if (r == 0 || !shared) bo = box_origin(rec, rs.o[r]);
            if (!__any_sync(kFull, mine)) continue;       // warp-uniform: no lane wants this ray slot tested
            if (COUNT) cnt.box_tests += mine;
            float tn, tf; Slabs sl; V3 ld;
            box_slabs(rec, bo, rs.d[r], tn, tf, sl, ld);
            const float tt = box_pick(tn, tf);
            // `mine` keeps the outcome a function of the ray alone (not of which lanes share its warp)
            if (mine && !box_reject(tn, tf) && tt > 0.0f) {
                if (tt < rs.best[r]) { rs.best[r] = tt; rs.bidx[r] = hot; }
                else if (!anyhit && tt == rs.best[r] && rs.bidx[r] >= 0 && hot_object(P, hot) < hot_object(P, rs.bidx[r])) rs.bidx[r] = hot;
            }
        }
    } else {
        // the discriminant of glsl:591, then the roots, ray slot by ray slot
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (r > 0 && !((want >> r) & 1u)) continue;       // warp-uniform
            if (r == 0 || !shared) so = sphere_origin(cc, rs.o[r]);
            const bool live = rs.best[r] >= live_min && ((want >> r) & 1u);
            if (COUNT) cnt.cull_tests += live;
            const float hb = __fmaf_rn(rs.d[r].x, so.oc.x, __fmaf_rn(rs.d[r].y, so.oc.y, __fmul_rn(rs.d[r].z, so.oc.z)));
            const float disc = __fmaf_rn(hb, hb, -__fmul_rn(rs.q[r].qa, so.qc));
            const bool go = live && disc >= 0.0f;
            if (!__any_sync(kFull, go)) continue;
            if (COUNT) cnt.sphere_tests += go;
            const float sq = sqrt_approx(fmaxf(disc, 0.0f));
            const float tn = __fmul_rn(-hb - sq, rs.q[r].iqa);
            const float tf = __fmul_rn(sq - hb, rs.q[r].iqa);
            const float tt = (disc < 0.0f) ? -1.0f : ((tn < 0.0f) ? tf : tn);
            if (go && tt > 0.0f) {
                if (tt < rs.best[r]) { rs.best[r] = tt; rs.bidx[r] = hot; }
                else if (!anyhit && tt == rs.best[r] && rs.bidx[r] >= 0 && hot_object(P, hot) < hot_object(P, rs.bidx[r])) rs.bidx[r] = hot;
            }
        }
    }
}

// packet_trace through the groups of rt_groups.cu (its own function: the object-by-object loop below keeps its
// registers and its place in the instruction cache).
template <bool COUNT, int NC>
__device__ __noinline__ void packet_trace_groups(const RtParams& P, WarpCtx& w, const Cone* __restrict__ cone_in, unsigned on,
                                                 float grow, float live_min, float pad1, float pad2, bool shared, const V3* __restrict__ o_in,
                                                 const V3* __restrict__ d_in, float* __restrict__ best_io, int* __restrict__ bidx_io,
                                                 RtCounters& cnt, unsigned want) {
    constexpr int R = kShadowBatch;
    RaySet rs;
    rayset_load(rs, o_in, d_in, best_io, bidx_io);
    Cone cone[NC];
#pragma unroll
    for (int k = 0; k < NC; ++k) cone[k] = cone_in[k];
    const int lane = w.lane;
    // Groups of 32 neighbouring objects (rt_groups.cu): one group sphere per lane and trip; the members of a
    // group that met a cone are one 512-byte read, one member per lane, tested as the loops below test them.
    // A group no cone meets has no member a cone meets (rt_cone.h: group_radius), so the same candidates
    // reach candidate_test; their order differs, which its accept rule does not see.
    const int ng = P.n_groups;
    float4 g_next = __ldg(P.grp + min(lane, ng - 1));
    for (int g0 = 0; g0 < ng; g0 += 32) {
        if (!__any_sync(kFull, rs.best[0] >= live_min || rs.best[1] >= live_min || rs.best[2] >= live_min)) break;
        const float4 gc = g_next;
        g_next = __ldg(P.grp + min(g0 + 32 + lane, ng - 1));
        unsigned cm = 0u;
#pragma unroll
        for (int k = 0; k < NC; ++k)
            if (NC == 1 || ((on >> k) & 1u)) cm |= cone_pass(cone[k], gc, group_radius(cone[k].apex, gc, grow)) ? (1u << k) : 0u;
        if (g0 + lane >= ng) cm = 0u;
        if (COUNT && g0 + lane < ng) cnt.packet_tests += (unsigned long long)(NC == 1 ? 1 : __popc(on));
        unsigned gmask = __ballot_sync(kFull, cm != 0u);
        if (gmask == 0u) continue;
        int gs = __ffs(gmask) - 1;
        gmask &= gmask - 1u;
        float4 c_n = __ldg(P.srec + (size_t)(g0 + gs) * RT_GROUP_SIZE + lane);
        uint32_t h_n = __ldg(P.sidx + (size_t)(g0 + gs) * RT_GROUP_SIZE + lane);
        unsigned cm_n = __shfl_sync(kFull, cm, gs);
        for (;;) {
            const float4 c = c_n;
            const int hot = (int)h_n;
            const unsigned cmg = cm_n;
            const bool more = gmask != 0u;                   // warp-uniform: the next group's records in flight
            if (more) {
                gs = __ffs(gmask) - 1;
                gmask &= gmask - 1u;
                c_n = __ldg(P.srec + (size_t)(g0 + gs) * RT_GROUP_SIZE + lane);
                h_n = __ldg(P.sidx + (size_t)(g0 + gs) * RT_GROUP_SIZE + lane);
                cm_n = __shfl_sync(kFull, cm, gs);
            }
            const float rp = packet_radius(c.w, hot >= P.n_box, grow, pad1, pad2);       // NaN: no object there
            bool pass = false;
#pragma unroll
            for (int k = 0; k < NC; ++k)
                if ((cmg >> k) & 1u) pass = pass | cone_pass(cone[k], c, rp);              // warp-uniform
            if (COUNT) cnt.packet_tests += (unsigned long long)(rp == rp ? __popc(cmg) : 0);
            unsigned mask = __ballot_sync(kFull, pass);
            while (mask) {
                const int src = __ffs(mask) - 1;
                mask &= mask - 1u;
                const int hot_s = __shfl_sync(kFull, hot, src);
                const float4 cc = make_float4(__shfl_sync(kFull, c.x, src), __shfl_sync(kFull, c.y, src),
                                              __shfl_sync(kFull, c.z, src), __shfl_sync(kFull, c.w, src));
                candidate_test<COUNT>(P, rs, hot_s, cc, shared, live_min, want, cnt);
            }
            if (!more) break;
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) { best_io[r] = rs.best[r]; bidx_io[r] = rs.bidx[r]; }
}

template <bool SMEM, bool COUNT, int NC>
__device__ __noinline__ void packet_trace_objects(const RtParams& P, WarpCtx& w, const Cone* __restrict__ cone_in, unsigned on,
                                                  float grow, float live_min, float pad1, float pad2, bool shared, const V3* __restrict__ o_in,
                                                  const V3* __restrict__ d_in, float* __restrict__ best_io, int* __restrict__ bidx_io,
                                                  RtCounters& cnt, unsigned want);
template <bool SMEM, bool COUNT, int NC>
__device__ __forceinline__ void packet_trace(const RtParams& P, WarpCtx& w, const Cone* __restrict__ cone_in, unsigned on,
                                             float grow, float live_min, float pad1, float pad2, bool shared, const V3* __restrict__ o_in,
                                             const V3* __restrict__ d_in, float* __restrict__ best_io, int* __restrict__ bidx_io,
                                             RtCounters& cnt, unsigned want = 7u) {
    if (P.n_groups > 0) packet_trace_groups<COUNT, NC>(P, w, cone_in, on, grow, live_min, pad1, pad2, shared, o_in, d_in, best_io, bidx_io, cnt, want);      // kernel-uniform
    else packet_trace_objects<SMEM, COUNT, NC>(P, w, cone_in, on, grow, live_min, pad1, pad2, shared, o_in, d_in, best_io, bidx_io, cnt, want);
}
// The whole object loop for one packet of 3 rays per lane: the primary rays of a work item (closest
// hit: live_min = smallest positive float, glsl:753-776) or shadow rays of the warp's hit points (any hit
// below t = 1: live_min = 1, glsl:816).  One copy in the instruction stream for both (out of line,
// runtime switches): with 16 warps per SM in different phases of a frame the large-scene kernels are
// instruction-cache bound, not issue bound.
template <bool SMEM, bool COUNT, int NC>
__device__ __noinline__ void packet_trace_objects(const RtParams& P, WarpCtx& w, const Cone* __restrict__ cone_in, unsigned on,
                                          float grow, float live_min, float pad1, float pad2, bool shared, const V3* __restrict__ o_in,
                                          const V3* __restrict__ d_in, float* __restrict__ best_io, int* __restrict__ bidx_io,
                                          RtCounters& cnt, unsigned want) {
    constexpr int R = kShadowBatch;
    RaySet rs;
    rayset_load(rs, o_in, d_in, best_io, bidx_io);
    Cone cone[NC];
#pragma unroll
    for (int k = 0; k < NC; ++k) cone[k] = cone_in[k];
    const bool anyhit = live_min >= 1.0f;
    const int lane = w.lane;
    // one cone: two objects per lane and iteration (64 objects per trip, two independent dependency
    // chains); three cones: one object per lane (the six chains of two would spill)
    constexpr int kHalves = NC == 1 ? 2 : 1;
    constexpr int W = 32 * kHalves;                              // objects per trip
    // Resident: the lists sit in shared memory.  Streamed: the warp's 6 KB of tile buffers are a ring of
    // trip-sized slots filled by per-lane 16-byte cp.async copies straight from L2, kSlots - 1 trips in
    // flight (a warp's pass over the scene is latency bound otherwise: one 3 KB tile ahead was not enough).
    constexpr int kSlots = (2 * kTileBytes) / (W * 16);
    constexpr int kAhead = kSlots - 1;
    for (int part = 0; part < 2; ++part) {
        // no live ray left in the warp (shadow rays: every one occluded, glsl:816 needs ANY occluder)
        if (!__any_sync(kFull, rs.best[0] >= live_min || rs.best[1] >= live_min || rs.best[2] >= live_min)) break;
        const bool sph = part != 0;
        const int n = sph ? P.n_sph : P.n_box;
        if (n == 0) continue;
        const int hot0 = sph ? P.n_box : 0;
        const float4* __restrict__ gp = sph ? P.sph_hot : P.box_cull;
        const uint32_t ra = SMEM ? smem_addr(smem_f4 + (sph ? w.sph_off4 : w.box_off4)) : w.buf_addr[0];
        auto issue = [&](int trip) {                              // this lane's records of one trip into its ring slot
            const uint32_t slot = ra + (uint32_t)(trip % kSlots) * (W * 16u);
#pragma unroll
            for (int hlf = 0; hlf < kHalves; ++hlf) {
                const int idx = min(trip * W + hlf * 32 + lane, n - 1);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(slot + (uint32_t)(hlf * 32 + lane) * 16u), "l"(gp + idx) : "memory");
            }
        };
        if (!SMEM) {
            __syncwarp();                                         // every lane is done with the ring (previous part / call)
#pragma unroll 1
            for (int t = 0; t < kAhead; ++t) {
                if (t * W < n) issue(t);
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
        }
        int trip = 0;
        for (int i0 = 0; i0 < n; i0 += W, ++trip) {
            // shadow packets: leave as soon as every ray of the warp is occluded (checked every 256 objects)
            if (anyhit && (i0 & 255) == 0 && i0 > 0 &&
                !__any_sync(kFull, rs.best[0] >= 1.0f || rs.best[1] >= 1.0f || rs.best[2] >= 1.0f)) break;
            uint32_t rt;                                          // shared-memory address of record i0
            if (SMEM) {
                rt = ra + (uint32_t)i0 * 16u;
            } else {
                __syncwarp();                                     // the slot of trip - 1 is free: it takes trip + kAhead
                if ((trip + kAhead) * W < n) issue(trip + kAhead);
                asm volatile("cp.async.commit_group;" ::: "memory");
                asm volatile("cp.async.wait_group %0;" ::"n"(kAhead) : "memory");
                __syncwarp();                                     // every lane's copies of this trip have landed
                rt = ra + (uint32_t)(trip % kSlots) * (W * 16u);
            }
            const int j0 = i0 + lane, j1 = j0 + 32;
            const int last = n - 1 - i0;                          // in-trip index of the list's last record
            const float4 c0 = lds128(rt + (uint32_t)min(lane, last) * 16u);
            const float rp0 = packet_radius(c0.w, sph, grow, pad1, pad2);
            float4 c1 = c0;
            float rp1 = rp0;
            if (kHalves == 2) {
                c1 = lds128(rt + (uint32_t)min(lane + 32, last) * 16u);
                rp1 = packet_radius(c1.w, sph, grow, pad1, pad2);
            }
            bool pass0 = false, pass1 = false;
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                if (NC == 1 || ((on >> k) & 1u)) {                  // no short circuit: lanes stay converged
                    pass0 = pass0 | cone_pass(cone[k], c0, rp0);
                    if (kHalves == 2) pass1 = pass1 | cone_pass(cone[k], c1, rp1);
                }
            }
            if (COUNT) cnt.packet_tests += (unsigned long long)((NC == 1 ? 1 : __popc(on)) * ((j0 < n) + (kHalves == 2 && j1 < n)));
            const unsigned mask0 = __ballot_sync(kFull, pass0 && j0 < n);
            const unsigned mask1 = kHalves == 2 ? __ballot_sync(kFull, pass1 && j1 < n) : 0u;
            if ((mask0 | mask1) == 0u) continue;
#pragma unroll 1
            for (int half = 0; half < kHalves; ++half) {
                unsigned mask = half ? mask1 : mask0;
                while (mask) {
                    const int k32 = 32 * half + __ffs(mask) - 1;  // in-trip index
                    mask &= mask - 1u;
                    const float4 cc = lds128(rt + (uint32_t)k32 * 16u);
                    candidate_test<COUNT>(P, rs, hot0 + i0 + k32, cc, shared, live_min, want, cnt);
                }
            }
        }
        if (!SMEM) {
            asm volatile("cp.async.wait_group 0;" ::: "memory");          // nothing of this pass may land later:
            __syncwarp();                                                 // the next user of the buffers may be a TMA tile
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) { best_io[r] = rs.best[r]; bidx_io[r] = rs.bidx[r]; }
}

// One tile, dispatched on its kind.
template <int R, bool ANYHIT, bool SHARED, bool COUNT, bool CULL>
__device__ __forceinline__ void any_tile(const RtParams& P, const float4* recs, int t,
                                         const V3 (&o)[kShadowBatch], const V3 (&d)[kShadowBatch],
                                         float (&best)[kShadowBatch], int (&bidx)[kShadowBatch], RtCounters& cnt) {
    const int nbt = n_box_tiles<CULL>(P);
    if (t < nbt) {
        const int first = t * box_tile_n<CULL>();
        const int n = min(box_tile_n<CULL>(), P.n_box - first);
        if (CULL) boxes_tile_cull<R, ANYHIT, SHARED, COUNT>(P, recs, n, first, o, d, best, bidx, cnt);
        else if (RT_STRICT) boxes_tile<R, ANYHIT, SHARED, COUNT>(recs, n, first, o, d, best, bidx, cnt.box_tests);
        else boxes_tile_pipe<R, ANYHIT, SHARED, COUNT>(recs, n, first, o, d, best, bidx, cnt.box_tests);
    } else {
        const int first = (t - nbt) * kTileSph;
        spheres_tile<R, ANYHIT, SHARED, COUNT, CULL>(P, recs, min(kTileSph, P.n_sph - first), first, o, d, best, bidx, cnt);
    }
}

// Small resident scene: both hot lists in one go, no tile bookkeeping.
template <int R, bool ANYHIT, bool SHARED, bool COUNT>
__device__ __forceinline__ void all_objects(const RtParams& P, const WarpCtx& w,
                                            const V3 (&o)[kShadowBatch], const V3 (&d)[kShadowBatch],
                                            float (&best)[kShadowBatch], int (&bidx)[kShadowBatch], RtCounters& cnt) {
    int first = 0;
#if !RT_STRICT
    if (ANYHIT && SHARED && P.room_skip) {                         // uniform: a property of the scene
        // The first box holds every lit light (the room around the built-in scene; host prologue, with margin).  A
        // shadow ray that starts inside it as well runs inside it all the way to its light: the exit lies beyond
        // t = 1 and the test cannot succeed (glsl:690-696: inside -> t = t_far; glsl:816 wants t < 1).  When that holds
        // for every live ray of the warp the box is left out -- shadow rays are any-hit, their order is free.
        // 0.99999: away from the last bits of the slab test's own rounding.
        const float4* __restrict__ rp = smem_f4 + w.box_off4;
        const float4 rec[RT_BOX_F4] = {rp[0], rp[1], rp[2]};
        const V3 so = box_origin(rec, o[0]);
        bool live = false;
#pragma unroll
        for (int r = 0; r < R; ++r) live = live || best[r] >= 1.0f;
        const bool in = fmaxf(fmaxf(fabsf(so.x), fabsf(so.y)), fabsf(so.z)) < 0.99999f;
        if (__all_sync(kFull, in || !live)) first = 1;
    }
#endif
    boxes_tile<R, ANYHIT, SHARED, COUNT, false>(smem_f4 + w.box_off4 + first * RT_BOX_F4, P.n_box - first, first, o, d, best, bidx, cnt.box_tests);
    spheres_tile<R, ANYHIT, SHARED, COUNT, false>(P, smem_f4 + w.sph_off4, P.n_sph, 0, o, d, best, bidx, cnt);
}

// get_closest_collision (glsl:738-782) for one ray per lane.
template <bool SMEM, bool REPACK, bool COUNT, bool CULL>
__device__ __forceinline__ void trace_closest(const RtParams& P, WarpCtx& w, V3 o0, V3 d0, float& best0, int& bidx0,
                                              RtCounters& cnt) {
    V3 o[kShadowBatch], d[kShadowBatch];
    float best[kShadowBatch]; int bidx[kShadowBatch];
    o[0] = o0; d[0] = d0; best[0] = best0; bidx[0] = -1;
    if (SMEM && !REPACK) {
        all_objects<1, false, false, COUNT>(P, w, o, d, best, bidx, cnt);
        best0 = best[0]; bidx0 = bidx[0];
        return;
    }
    const int n_tiles = n_box_tiles<CULL>(P) + n_sph_tiles(P);
    tile_begin<SMEM, CULL>(P, w);
    for (int t = 0; t < n_tiles; ++t) {
        const float4* recs = tile_acquire<SMEM, CULL>(P, w, t, n_tiles);
        any_tile<1, false, false, COUNT, CULL>(P, recs, t, o, d, best, bidx, cnt);
    }
    best0 = best[0]; bidx0 = bidx[0];
}

// -------------------------------------------------------------------------
// The shadow rays of one hit point per lane (glsl:807-819): common origin so,
// directions sd[r] (light - p, un-normalised), sb[r] = 1 for a live slot, 0 for
// an unused one.  On return sb[r] < 1 iff the ray is occluded.
//
// Starts as 3 rays per lane with the shared origin.  Rays drop out as soon as
// ANY object occludes them, so at every tile boundary the warp counts its live
// rays and, when they fit in fewer slots, re-packs them densely across the 32
// lanes through a per-warp shared-memory queue (then 2, then 1 ray per lane, each
// with its own origin).  Dead rays therefore stop costing issue slots, and a warp
// whose rays are all occluded leaves the object loop.
// -------------------------------------------------------------------------
template <bool SMEM, bool REPACK, bool COUNT, bool CULL, bool LREC>
__device__ __forceinline__ void trace_shadow(const RtParams& P, WarpCtx& w, V3 so, const V3 (&sd)[kShadowBatch],
                                             float (&sb)[kShadowBatch], int n_slots, RtCounters& cnt) {
    V3 o[kShadowBatch], d[kShadowBatch];
    float best[kShadowBatch]; int bidx[kShadowBatch]; int src[kShadowBatch];
#pragma unroll
    for (int r = 0; r < kShadowBatch; ++r) { o[r] = so; d[r] = sd[r]; best[r] = sb[r]; bidx[r] = -1; src[r] = w.lane * kShadowBatch + r; }
    int r_cur = n_slots;           // slots in use: 1..3 with the shared origin, then 2 or 1 after a re-pack
    bool shared = true;
    unsigned own_occ = 0u;
    const unsigned lt_mask = (1u << w.lane) - 1u;
    const int n_tiles = LREC ? n_lrec_tiles(P) : n_box_tiles<CULL>(P) + n_sph_tiles(P);
    bool started = false;

    if (!REPACK) {
        // A handful of objects (the built-in scene has 5): the loop is shorter than one re-pack.
        // Chosen at launch (kRepackMinObjects); keeps the small-scene kernel small (instruction cache).
        if (SMEM) {
            // n_slots (1..3 rays per lane, warp-uniform) picks the loop: no issue slots for unused rays
            if (n_slots == 3) all_objects<3, true, true, COUNT>(P, w, o, d, best, bidx, cnt);
            else if (n_slots == 2) all_objects<2, true, true, COUNT>(P, w, o, d, best, bidx, cnt);
            else all_objects<1, true, true, COUNT>(P, w, o, d, best, bidx, cnt);
        } else {
            tile_begin<SMEM, CULL>(P, w);
            for (int t = 0; t < n_tiles; ++t) {
                const float4* recs = tile_acquire<SMEM, CULL>(P, w, t, n_tiles);
                any_tile<kShadowBatch, true, true, COUNT, CULL>(P, recs, t, o, d, best, bidx, cnt);
            }
        }
#pragma unroll
        for (int r = 0; r < kShadowBatch; ++r) sb[r] = best[r];
        return;
    }

    for (int t = 0; t < n_tiles; ++t) {
        // live rays of the warp, slot by slot
        unsigned m[kShadowBatch];
        int live = 0;
#pragma unroll
        for (int r = 0; r < kShadowBatch; ++r) {
            m[r] = (r < r_cur) ? __ballot_sync(kFull, best[r] >= 1.0f) : 0u;
            live += __popc(m[r]);
        }
        if (live == 0) {
            if (started) tile_drain<SMEM>(w, t - 1, n_tiles);   // tile t is in flight
            break;
        }
        const int new_r = (live + 31) >> 5;
        if (new_r < r_cur) {
            // ---- re-pack the live rays densely: queue entry = (origin, direction, source id) ----
            if (shared) {
#pragma unroll
                for (int r = 0; r < kShadowBatch; ++r) {
                    if (sb[r] >= 1.0f && best[r] < 1.0f) own_occ |= 1u << r;
                    w.occ[w.lane * kShadowBatch + r] = 0;
                }
            } else {
#pragma unroll
                for (int r = 0; r < kShadowBatch; ++r)
                    if (r < r_cur && src[r] >= 0 && best[r] < 1.0f) w.occ[src[r]] = 1;
            }
            int base = 0;
#pragma unroll
            for (int r = 0; r < kShadowBatch; ++r) {
                if (r < r_cur) {
                    if (best[r] >= 1.0f) {
                        float* q = w.queue + (base + __popc(m[r] & lt_mask)) * 8;
                        const V3 oo = shared ? o[0] : o[r];
                        *reinterpret_cast<float4*>(q) = make_float4(oo.x, oo.y, oo.z, d[r].x);
                        *reinterpret_cast<float4*>(q + 4) = make_float4(d[r].y, d[r].z, __int_as_float(src[r]), 0.0f);
                    }
                    base += __popc(m[r]);
                }
            }
            __syncwarp();
#pragma unroll
            for (int r = 0; r < kShadowBatch; ++r) {
                const int idx = r * 32 + w.lane;
                if (r < new_r && idx < live) {
                    const float4 a = *reinterpret_cast<const float4*>(w.queue + idx * 8);
                    const float4 b = *reinterpret_cast<const float4*>(w.queue + idx * 8 + 4);
                    o[r] = mk(a.x, a.y, a.z); d[r] = mk(a.w, b.x, b.y); src[r] = __float_as_int(b.z);
                    best[r] = 1.0f;
                } else {
                    o[r] = mk(0.0f, 0.0f, 0.0f); d[r] = mk(0.0f, 0.0f, 1.0f); src[r] = -1;
                    best[r] = 0.0f;
                }
            }
            __syncwarp();
            r_cur = new_r;
            shared = false;
        }
        if (LREC) {
            // per-light records: no origin work in the loop, so "shared" only matters to the re-pack above
            if (!started) { if (!SMEM && w.lane == 0) lrec_issue(P, w, 0); started = true; }
            const float4* lrecs = lrec_acquire<SMEM>(P, w, t, n_tiles);
            if (r_cur == 3) lights_tile<3, COUNT>(P, lrecs, t, o, d, best, src, cnt);
            else if (r_cur == 2) lights_tile<2, COUNT>(P, lrecs, t, o, d, best, src, cnt);
            else lights_tile<1, COUNT>(P, lrecs, t, o, d, best, src, cnt);
            continue;
        }
        if (!started) { tile_begin<SMEM, CULL>(P, w); started = true; }
        const float4* recs = tile_acquire<SMEM, CULL>(P, w, t, n_tiles);
        if (shared) {
            if (r_cur == 3) any_tile<3, true, true, COUNT, CULL>(P, recs, t, o, d, best, bidx, cnt);
            else if (r_cur == 2) any_tile<2, true, true, COUNT, CULL>(P, recs, t, o, d, best, bidx, cnt);
            else any_tile<1, true, true, COUNT, CULL>(P, recs, t, o, d, best, bidx, cnt);
        } else if (r_cur == 2) any_tile<2, true, false, COUNT, CULL>(P, recs, t, o, d, best, bidx, cnt);
        else any_tile<1, true, false, COUNT, CULL>(P, recs, t, o, d, best, bidx, cnt);
    }

    if (shared) {
#pragma unroll
        for (int r = 0; r < kShadowBatch; ++r) sb[r] = best[r];
    } else {
        // holders report, owners collect
#pragma unroll
        for (int r = 0; r < kShadowBatch; ++r)
            if (r < r_cur && src[r] >= 0 && best[r] < 1.0f) w.occ[src[r]] = 1;
        __syncwarp();
#pragma unroll
        for (int r = 0; r < kShadowBatch; ++r) {
            const bool occluded = ((own_occ >> r) & 1u) || w.occ[w.lane * kShadowBatch + r] != 0;
            if (sb[r] >= 1.0f) sb[r] = occluded ? 0.5f : 1.0f;
        }
        __syncwarp();
    }
}

// -------------------------------------------------------------------------
// Shadow rays as packets: one cone per light slot, apex at the light, around the directions from
// the light to the warp's hit points.  A shadow ray runs from p + 0.01 n along L - p (glsl:808-809),
// so its line passes within 0.01 of L; the bounding radii grow by 0.0101 (`grow`) to cover that, and
// nothing further from L than the farthest hit point (+ 0.0101) can occlude (t < 1, glsl:816).
// Returns false (nothing traced) when some slot's directions spread beyond kConeMinCos.
// -------------------------------------------------------------------------
__device__ __forceinline__ bool shadow_cones(const WarpCtx& w, const V3 (&lp)[kShadowBatch], const V3 (&sd)[kShadowBatch],
                                             const float (&sb)[kShadowBatch], Cone (&cone)[kShadowBatch], unsigned& on) {
    on = 0u;
    bool ok = true;
#pragma unroll
    for (int k = 0; k < kShadowBatch; ++k) {
        const bool act = sb[k] >= 1.0f;
        const unsigned am = __ballot_sync(kFull, act);
        if (am == 0u) continue;
        on |= 1u << k;
        const float l2 = __fmaf_rn(sd[k].x, sd[k].x, __fmaf_rn(sd[k].y, sd[k].y, __fmul_rn(sd[k].z, sd[k].z)));
        const float il = rsqrtf(l2);
        const V3 u = mk(-sd[k].x * il, -sd[k].y * il, -sd[k].z * il);      // from the light toward the hit point
        // axis: a lane near the middle of the 8x4 tile if it is active, else the first active lane
        const unsigned mid = am & 0x00181800u;
        const int lead = __ffs(mid ? mid : am) - 1;
        cone[k].apex = lp[k];
        cone[k].axis = mk(__shfl_sync(kFull, u.x, lead), __shfl_sync(kFull, u.y, lead), __shfl_sync(kFull, u.z, lead));
        const V3 x = cross3(cone[k].axis, u);
        const float s2 = __fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z)));
        const float cs = dot3(cone[k].axis, u);
        ok = ok && __all_sync(kFull, !act || cs >= kConeMinCos);             // NaN (zero-length ray) -> false
        const float s2max = warp_max_nonneg(act ? fminf(s2, 1.0f) : 0.0f);
        const float l2max = warp_max_nonneg(act ? l2 : 0.0f);
        cone_close(cone[k], s2max, __fmaf_rn(sqrt_approx(l2max), 1.0001f, 0.0101f));
    }
    return ok;
}

// -------------------------------------------------------------------------
// Secondary rays (the reflection / refraction children of glsl:988-1027) as packets.  One ray per lane, origins
// and directions differ from lane to lane, but the children of neighbouring pixels that hit the same face
// leave almost parallel from almost the same place.  Lanes are grouped greedily around a leader (direction
// within P.group_cos of the leader's, origin within sqrt(rg2) of it); a group is traced as ONE packet: a cone
// from the groups's mean origin around its mean direction with every bounding radius grown by the distance of
// the farthest group origin from that apex (a ray shifted to the apex moves by at most that much).  Still every
// object for every group, in order -- 64 objects per trip instead of one per ray.  At most kMaxGroups groups;
// lanes left over are returned (true) for the per-ray loops.  A ray with a zero or NaN direction (refract()'s
// total internal reflection, glsl:1023) misses everything in the per-ray loops too and is dropped here.
// -------------------------------------------------------------------------
constexpr int kMaxGroups = 12;
template <bool SMEM, bool COUNT>
__device__ __forceinline__ bool secondary_packets(const RtParams& P, WarpCtx& w, bool act, V3 o, V3 d, float rg2, float pad1, float pad2,
                                                  float& best, int& bidx, RtCounters& cnt) {
    const float l2 = dot3(d, d);
    bool todo = act && l2 > 0.0f && l2 < 1e30f;
    const V3 u = d * rsqrtf(l2);
    for (int round = 0; round < kMaxGroups; ++round) {
        const unsigned rem = __ballot_sync(kFull, todo);
        if (rem == 0u) break;
        const int lead = __ffs(rem) - 1;
        const V3 lu = mk(__shfl_sync(kFull, u.x, lead), __shfl_sync(kFull, u.y, lead), __shfl_sync(kFull, u.z, lead));
        const V3 lo = mk(__shfl_sync(kFull, o.x, lead), __shfl_sync(kFull, o.y, lead), __shfl_sync(kFull, o.z, lead));
        const V3 dl = o - lo;
        const bool in_g = todo && (w.lane == lead || (dot3(u, lu) >= P.group_cos && dot3(dl, dl) <= rg2));
        float acc[7] = {in_g ? u.x : 0.0f, in_g ? u.y : 0.0f, in_g ? u.z : 0.0f, in_g ? o.x : 0.0f, in_g ? o.y : 0.0f, in_g ? o.z : 0.0f, in_g ? 1.0f : 0.0f};
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
            for (int c = 0; c < 7; ++c) acc[c] += __shfl_xor_sync(kFull, acc[c], off);
        }
        Cone gc[1];
        const float inv_n = 1.0f / acc[6];
        gc[0].apex = mk(acc[3] * inv_n, acc[4] * inv_n, acc[5] * inv_n);
        gc[0].axis = normalize3(mk(acc[0], acc[1], acc[2]));
        const V3 x = cross3(gc[0].axis, u);
        const V3 da = o - gc[0].apex;
        const float s2 = warp_max_nonneg(in_g ? fminf(__fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z))), 1.0f) : 0.0f);
        const float ro2 = warp_max_nonneg(in_g ? dot3(da, da) : 0.0f);
        cone_close(gc[0], s2, 1e18f);
        const float grow = __fmaf_rn(sqrt_approx(ro2), 1.001f, pad1);
        V3 o3[kShadowBatch] = {o, o, o}, d3[kShadowBatch] = {d, d, d};
        float b3[kShadowBatch] = {in_g ? kFar : 0.0f, 0.0f, 0.0f};
        int i3[kShadowBatch] = {-1, -1, -1};
        packet_trace<SMEM, COUNT, 1>(P, w, gc, 1u, grow, 1e-45f, pad1, pad2, false, o3, d3, b3, i3, cnt, 1u);
        if (COUNT && w.lane == 0) cnt.group_passes++;
        if (in_g) { best = b3[0]; bidx = i3[0]; todo = false; }
    }
    return todo;
}

// -------------------------------------------------------------------------
// Closest-hit record: everything ads_phong_lighting and the bounce phases need
// (glsl:561-573 Collision + the object's material).
// -------------------------------------------------------------------------
struct Hit {
    V3 p, n;
    bool inside;
    int obj;
};

// One row of mat4 * vec4(v, 1) / mat4 * vec4(v, 0) with the shader's association ((m0 x + m1 y) + m2 z) + m3 w.
__device__ __forceinline__ float row_pt(const float* __restrict__ m, V3 v) {
    return __fadd_rn(madd(m[2], v.z, madd(m[1], v.y, __fmul_rn(m[0], v.x))), m[3]);
}
__device__ __forceinline__ float row_dir(const float* __restrict__ m, V3 v) {
#if RT_STRICT
    return __fadd_rn(madd(m[2], v.z, madd(m[1], v.y, __fmul_rn(m[0], v.x))), __fmul_rn(m[3], 0.0f));
#else
    return madd(m[2], v.z, madd(m[1], v.y, __fmul_rn(m[0], v.x)));
#endif
}
// Re-evaluates the winning primitive to produce p, n, inside (glsl:627-637 sphere,
// glsl:686-721 box).  t is the value the trace loop accepted.
template <bool SMEM, bool CULL>
__device__ __forceinline__ Hit resolve_hit(const RtParams& P, const WarpCtx& w, V3 o, V3 d, float t, int bidx) {
    // the winner's hot record: shared memory when resident, global/L2 when streaming (once per ray)
    const float4* box_hot = (SMEM && !CULL) ? smem_f4 + w.box_off4 : P.box_hot;
    const float4* sph_hot = SMEM ? smem_f4 + w.sph_off4 : P.sph_hot;
    Hit h;
    if (bidx >= P.n_box) {
        const int j = bidx - P.n_box;
        const float4 c = sph_hot[j];
        bool inside;
        (void)sphere_t(sphere_origin(c, o), d, ray_q(d), inside);
        h.obj = __ldg(P.sph_obj + j);
        h.inside = inside;
        h.p = axpy3(d, t, o);                                       // glsl:628
        h.n = normalize3(mk(h.p.x - c.x, h.p.y - c.y, h.p.z - c.z));  // glsl:631
        if (inside) h.n = h.n * -1.0f;                              // glsl:634-637
        return h;
    }
    h.obj = __ldg(P.box_obj + bidx);
    const RtCold* __restrict__ c = P.cold + h.obj;
    float4 rec[RT_BOX_F4];
#pragma unroll
    for (int k = 0; k < RT_BOX_F4; ++k) rec[k] = box_hot[bidx * RT_BOX_F4 + k];
    float tn, tf; Slabs s; V3 ld;
    box_slabs(rec, box_origin(rec, o), d, tn, tf, s, ld);
    h.inside = tn < 0.0f;                                           // glsl:690-696
    const float by = h.inside ? s.t2y : s.t1y, bz = h.inside ? s.t2z : s.t1z;
    int face = 0;                                                   // glsl:699-704
    if (t == by) face = 1;
    else if (t == bz) face = 2;
    const float* m = c->nmat;                                       // glsl:718, column-major
#if RT_STRICT
    // ray in the box's (unscaled) local space, as the shader computes it (glsl:655-656)
    const float* m4 = c->w2l;
    V3 rs, rd;
    rs = mk(row_pt(m4, o), row_pt(m4 + 4, o), row_pt(m4 + 8, o));
    rd = mk(row_dir(m4, d), row_dir(m4 + 4, d), row_dir(m4 + 8, d));
    const float rdf = face == 0 ? rd.x : (face == 1 ? rd.y : rd.z);
    V3 nl = mk(face == 0 ? 1.0f : 0.0f, face == 1 ? 1.0f : 0.0f, face == 2 ? 1.0f : 0.0f);   // glsl:707-708
    if (rdf > 0.0f) nl = nl * -1.0f;                                // glsl:712-715
    h.n.x = madd(m[6], nl.z, madd(m[3], nl.y, __fmul_rn(m[0], nl.x)));
    h.n.y = madd(m[7], nl.z, madd(m[4], nl.y, __fmul_rn(m[1], nl.x)));
    h.n.z = madd(m[8], nl.z, madd(m[5], nl.y, __fmul_rn(m[2], nl.x)));
    const V3 ql = axpy3(rd, t, rs);                                 // glsl:721
    const float* l = c->l2w;
    h.p = mk(row_pt(l, ql), row_pt(l + 4, ql), row_pt(l + 8, ql));
#else
    // Shipped kernel: the same quantities without the detour through the box's local space.
    //  * normal = +-column `face` of the normal matrix (glsl:707-718: the matrix times a signed unit vector);
    //    the sign is the sign of the local direction component (glsl:712), which the unit-frame direction
    //    `ld` of the slab test shares (its rows are the local rows scaled by 1 / half-extent > 0);
    //  * hit point = o + t d, which is what local_to_world * (ray_start + t * ray_dir) (glsl:721) equals before
    //    rounding (ulp-level difference, stated in DESIGN.md 7).
    const float ldf = face == 0 ? ld.x : (face == 1 ? ld.y : ld.z);
    const float sg = ldf > 0.0f ? -1.0f : 1.0f;
    const float* mc = m + 3 * face;
    h.n = mk(__fmul_rn(mc[0], sg), __fmul_rn(mc[1], sg), __fmul_rn(mc[2], sg));
    h.p = axpy3(d, t, o);
#endif
    return h;
}

// Jitter extension (identical integer hash to the oracle; spp == 1 -> (0,0) exactly).
__device__ __forceinline__ uint32_t mix32(uint32_t h) {
    h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
    return h;
}
__device__ __forceinline__ void jitter(const RtParams& P, int x, int y, int s, float& jx, float& jy) {
    if (P.spp <= 1) { jx = 0.0f; jy = 0.0f; return; }
    uint32_t h = P.seed;
    h = mix32(h ^ ((uint32_t)x * 0x9E3779B1u));
    h = mix32(h ^ ((uint32_t)y * 0x85EBCA77u));
    h = mix32(h ^ ((uint32_t)s * 0xC2B2AE3Du));
    uint32_t h2 = mix32(h + 0x68E31DA4u);
    float ux = (float)(h >> 8) * (1.0f / 16777216.0f);
    float uy = (float)(h2 >> 8) * (1.0f / 16777216.0f);
    const int n = P.spp_n;
    if (n > 0) {
        int cx = s % n, cy = s / n;
        jx = ((float)cx + ux) / (float)n;
        jy = ((float)cy + uy) / (float)n;
    } else {
        jx = ux; jy = uy;
    }
}

// glsl:377-392 with inverse(proj*view) hoisted to the host prologue.
__device__ __forceinline__ V3 primary_dir(const RtParams& P, const RtCameraDev* __restrict__ cam,
                                          int px, int py, float jx, float jy) {
    const float vx = ((float)(px - P.width / 2) + jx) / (float)(P.width / 2);
    const float vy = ((float)(py - P.height / 2) + jy) / (float)(P.height / 2);
    const float* m = cam->inv_pv;   // column-major: element (row r, col c) = m[4*c + r]
    float s[4], e[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const float xy = madd(m[4 + r], vy, __fmul_rn(m[r], vx));                  // (m0 vx + m1 vy)
        s[r] = __fadd_rn(madd(m[8 + r], 0.5f, xy), m[12 + r]);                      // ... + m2 * 0.5 + m3 * 1
        e[r] = __fadd_rn(__fadd_rn(xy, m[8 + r]), m[12 + r]);                       // ... + m2 * 1 + m3 * 1
    }
    V3 ws = mk(fdiv(s[0], s[3]), fdiv(s[1], s[3]), fdiv(s[2], s[3]));
    V3 we = mk(fdiv(e[0], e[3]), fdiv(e[1], e[3]), fdiv(e[2], e[3]));
    return normalize3(we - ws);
}

__device__ __forceinline__ uint32_t unorm8(float f) {
    // imageStore to RGBA8 (glsl:404, main.cpp:223): clamp, scale, round-to-nearest-even; NaN -> 0
    if (f != f) return 0u;
    float c = fminf(fmaxf(f, 0.0f), 1.0f);
    return (uint32_t)__float2int_rn(c * 255.0f);
}

// -------------------------------------------------------------------------
// One sample: recursive_raytrace (glsl:1071-1105) as a weighted pending-ray stack.
// All 32 lanes run the loop together; a lane without work rides along inactive.
// -------------------------------------------------------------------------
// -------------------------------------------------------------------------
// ads_phong_lighting's light loop (glsl:797-834) for a small scene whose 1..3 lit lights ride in the kernel
// parameters: NS shadow rays with the shared origin through all_objects, then the diffuse / specular terms of the
// lights that are not occluded -- straight-line code, light data as constant-bank operands.  Same arithmetic as the
// general path in trace_sample.
// -------------------------------------------------------------------------
template <int NS, bool COUNT>
__device__ __forceinline__ void shade_small(const RtParams& P, const WarpCtx& w, const Hit& h, const RtCold* __restrict__ mat,
                                            V3 view, bool shade, float (&dif)[4], float (&spc)[4], RtCounters& cnt) {
    V3 o[kShadowBatch], d[kShadowBatch];
    float best[kShadowBatch]; int bidx[kShadowBatch];
    const V3 so = shade ? axpy3(h.n, 0.01f, h.p) : mk(0.0f, 0.0f, 0.0f);                    // glsl:808
#pragma unroll
    for (int r = 0; r < kShadowBatch; ++r) {
        o[r] = so; bidx[r] = -1;
        if (r < NS) {
            const float4 lp = P.lit_pos[r];
            d[r] = shade ? mk(lp.x - h.p.x, lp.y - h.p.y, lp.z - h.p.z) : mk(0.0f, 0.0f, 1.0f);   // glsl:809
            best[r] = shade ? 1.0f : 0.0f;                                                   // glsl:816: shadowed iff closest t < 1
        } else {
            d[r] = mk(0.0f, 0.0f, 1.0f); best[r] = 0.0f;
        }
    }
    all_objects<NS, true, true, COUNT>(P, w, o, d, best, bidx, cnt);
    const float4 md = *reinterpret_cast<const float4*>(mat->diffuse), ms = *reinterpret_cast<const float4*>(mat->specular);
    const float shin = mat->shininess;
#pragma unroll
    for (int r = 0; r < NS; ++r) {
        if (shade && !(best[r] < 1.0f)) {
            if (COUNT) cnt.shaded_lights++;
            const V3 light_dir = normalize3(d[r]);                                           // glsl:804
            const V3 light_ref = normalize3(reflect3(-light_dir, h.n));                      // glsl:825
            const float cos_theta = dot3(light_dir, h.n);                                    // glsl:826
            const float cos_phi = dot3(view, light_ref);                                     // glsl:827
            const float kd = fmaxf(cos_theta, 0.0f);
            const float ks = fpow(fmaxf(cos_phi, 0.0f), shin);                               // glsl:832
            const float4 ld4 = P.lit_dif[r], ls4 = P.lit_spc[r];
            dif[0] += ld4.x * md.x * kd; dif[1] += ld4.y * md.y * kd; dif[2] += ld4.z * md.z * kd; dif[3] += ld4.w * md.w * kd;   // glsl:830
            spc[0] += ls4.x * ms.x * ks; spc[1] += ls4.y * ms.y * ks; spc[2] += ls4.z * ms.z * ks; spc[3] += ls4.w * ms.w * ks;   // glsl:832
        }
    }
}

// -------------------------------------------------------------------------
// Phong terms with explicit roundings (madd / __fmul_rn / __fadd_rn): trace_sample and trace_sample3 evaluate the
// same pixels along different code paths and must agree bit for bit, so nothing here is left to the compiler's
// FMA contraction (which is chosen per inlined instance).
// -------------------------------------------------------------------------
// diffuse += L.diffuse * M.diffuse * max(n.l, 0); specular += L.specular * M.specular * pow(max(v.r, 0), shininess)  (glsl:822-833)
__device__ __forceinline__ void phong_light(const RtLightDev* __restrict__ L, const RtCold* __restrict__ mat, V3 sd, V3 n, V3 view,
                                            float (&dif)[4], float (&spc)[4]) {
    const V3 light_dir = normalize3(sd);                                     // glsl:804
    const V3 light_ref = normalize3(reflect3(-light_dir, n));                // glsl:825
    const float kd = fmaxf(dot3(light_dir, n), 0.0f);                        // glsl:826
    const float ks = fpow(fmaxf(dot3(view, light_ref), 0.0f), mat->shininess);   // glsl:827, 832
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        dif[k] = madd(__fmul_rn(L->diffuse[k], mat->diffuse[k]), kd, dif[k]);        // glsl:830
        spc[k] = madd(__fmul_rn(L->specular[k], mat->specular[k]), ks, spc[k]);      // glsl:832
    }
}
// (ambient + diffuse + specular + emissive).rgb * .a  (glsl:837-839), added to the pixel with the node's weight
__device__ __forceinline__ V3 phong_accumulate(V3 color, float wp, const float (&amb)[4], const float (&dif)[4], const float (&spc)[4],
                                               const RtCold* __restrict__ mat) {
    float pc[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) pc[k] = __fadd_rn(__fadd_rn(__fadd_rn(amb[k], dif[k]), spc[k]), mat->emissive[k]);
    return mk(madd(wp, __fmul_rn(pc[0], pc[3]), color.x), madd(wp, __fmul_rn(pc[1], pc[3]), color.y), madd(wp, __fmul_rn(pc[2], pc[3]), color.z));
}

// Pending rays of one lane: origin, direction, weight, remaining depth.  BOUNCE = false (MAX_RAYTRACE_DEPTH = 0,
// the shader as shipped, glsl:22): the tree is its root, there is nothing to keep -- no local-memory frame at all.
template <bool BOUNCE> struct RayStack;
template <> struct RayStack<true> {
    float4 stk[kStack][2];          // two 16-byte local-memory accesses per push / pop
    int sp;
    __device__ __forceinline__ void init(bool valid, V3 o, V3 d, int depth) { sp = 0; if (valid) push(o, d, 1.0f, depth); }
    __device__ __forceinline__ bool pending() const { return sp > 0; }
    __device__ __forceinline__ void clear() { sp = 0; }
    __device__ __forceinline__ void push(V3 o, V3 d, float wgt, int depth) {
        stk[sp][0] = make_float4(o.x, o.y, o.z, d.x);
        stk[sp][1] = make_float4(d.y, d.z, wgt, __int_as_float(depth));
        ++sp;
    }
    __device__ __forceinline__ void pop(V3& o, V3& d, float& wgt, int& depth) {
        --sp;
        const float4 a = stk[sp][0], b = stk[sp][1];
        o = mk(a.x, a.y, a.z);
        d = mk(a.w, b.x, b.y);
        wgt = b.z; depth = __float_as_int(b.w);
    }
};
template <> struct RayStack<false> {
    V3 o0, d0;
    bool live;
    __device__ __forceinline__ void init(bool valid, V3 o, V3 d, int) { o0 = o; d0 = d; live = valid; }
    __device__ __forceinline__ bool pending() const { return live; }
    __device__ __forceinline__ void clear() { live = false; }
    __device__ __forceinline__ void push(V3, V3, float, int) {}
    __device__ __forceinline__ void pop(V3& o, V3& d, float& wgt, int& depth) { o = o0; d = d0; wgt = 1.0f; depth = 0; live = false; }
};

template <bool SMEM, bool REPACK, bool COUNT, bool CULL, bool BOUNCE = true>
__device__ __forceinline__ V3 trace_sample(const RtParams& P, WarpCtx& w, bool valid, V3 org, V3 dir,
                                           bool have_first, float first_t, int first_idx, bool have_sb, const float (&first_sb)[kShadowBatch],
                                           float pk_pad1, float pk_pad2, RtCounters& cnt) {
    RayStack<BOUNCE> stack;
    stack.init(valid, org, dir, P.max_depth);
    V3 color = mk(0.0f, 0.0f, 0.0f);
    int iters = 0;   // loop iterations the shader's stack machine would have spent (glsl:1081-1094)

    while (__any_sync(kFull, stack.pending())) {
        const bool act = stack.pending();
        V3 o[1], d[1];
        float wgt = 0.0f; int depth = 0;
        float best[1]; int bidx[1];
        bidx[0] = -1;
        if (act) {
            stack.pop(o[0], d[0], wgt, depth);
            if (!BOUNCE) depth = 0;
            best[0] = kFar;
        } else {
            o[0] = mk(0.0f, 0.0f, 0.0f); d[0] = mk(0.0f, 0.0f, 1.0f);
            best[0] = 0.0f;
        }
        // PHASE 0 (glsl:956-967)
        bool root_sb = false;       // root node whose shadow rays the caller has traced (root_shadows)
        const bool first_now = have_first;
        if (have_first) {           // warp-uniform: the root ray was traced by the caller
            if (act) { best[0] = first_t; bidx[0] = first_idx; }
            have_first = false;
            root_sb = have_sb;
        }
        if (!RT_STRICT && REPACK && CULL) {
            bool todo = act && !first_now;      // lanes whose ray goes through the per-ray loops
            if (!first_now && P.packet && P.group_k > 0.0f)
                todo = secondary_packets<SMEM, COUNT>(P, w, act, o[0], d[0], pk_pad2 * 2.5e5f * P.group_k, pk_pad1, pk_pad2, best[0], bidx[0], cnt);
            if (__any_sync(kFull, todo)) {
                float b1 = todo ? kFar : 0.0f; int i1 = -1;
                if (COUNT && w.lane == 0) cnt.loop_passes++;
                trace_closest<SMEM, REPACK, COUNT, CULL>(P, w, o[0], d[0], b1, i1, cnt);
                if (todo) { best[0] = b1; bidx[0] = i1; }
            }
        } else if (!first_now) {
            trace_closest<SMEM, REPACK, COUNT, CULL>(P, w, o[0], d[0], best[0], bidx[0], cnt);
        }
        const bool hit = act && bidx[0] >= 0;
        if (COUNT && act) { cnt.rays++; cnt.nodes++; cnt.hit_nodes += hit; }
        if (COUNT) cnt.warp_iters++;
        if (act) iters += hit ? 6 : 1;

        Hit h;
        const RtCold* __restrict__ mat = P.cold;
        if (hit) {
            h = resolve_hit<SMEM, CULL>(P, w, o[0], d[0], best[0], bidx[0]);
            mat = P.cold + h.obj;
        }
        // PHASE 1: ads_phong_lighting (glsl:789-840), shadow rays kShadowBatch at a time
        float amb[4] = {0, 0, 0, 0}, dif[4] = {0, 0, 0, 0}, spc[4] = {0, 0, 0, 0};
        V3 view = mk(0.0f, 0.0f, 0.0f);
        // Bounce weights first (glsl:979-1054): final = mix(mix(phong, refl, kr), refr, kt) is linear in
        // its three colours, so phong enters the pixel with the product weight wp.
        float kr = 0.0f, kt = 0.0f, wp = 0.0f, wr = 0.0f;
        bool reflected = false, refracted = false;
        if (hit) {
            kr = mat->reflectivity; kt = mat->transparency;
            reflected = BOUNCE && depth > 0 && kr > 0.0f;
            refracted = BOUNCE && depth > 0 && kt > 0.0f;
            wp = wgt; wr = wgt * kr;
            if (reflected) wp = wp * (1.0f - kr);
            if (refracted) { wp = wp * (1.0f - kt); wr = wr * (1.0f - kt); }
        }
        // Dead-work elimination with identical results: a surface whose Phong colour enters with
        // weight exactly 0 (a perfect mirror above depth 0: mix(phong, refl, 1.0)) needs neither
        // lighting nor shadow rays; a light with all-zero diffuse and specular (the shader's "world
        // ambient" light 0, glsl:202-207) adds 0 whether shadowed or not, so its shadow ray is not
        // traversed.  The reference casts these rays; they are still COUNTED as rays (cnt.rays is the
        // reference's ray count) and reported separately as not traversed.
        const bool shade = hit && wp != 0.0f;
        if (COUNT && hit && !shade) { cnt.rays += P.n_lights; cnt.shadow_rays += P.n_lights; cnt.pad += P.n_lights; }
        if (shade) view = normalize3(-d[0]);                        // glsl:827 normalize(-r.dir)
        if (__any_sync(kFull, shade)) {
            // P.n_lit_batch >= 0 (fast kernel, at most 3 lit lights): the host prologue has summed the
            // lights' ambient terms (glsl:801) and listed the lit lights, so there is one shadow batch and
            // no per-node loop over the light list.  Otherwise (strict kernel, > 3 lit lights): the loops.
            const bool one_batch = !RT_STRICT && P.n_lit_batch >= 0;
            if (!RT_STRICT && !REPACK && one_batch) {
                // small scenes: one straight-line batch, light data from the kernel parameters
                if (shade) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) amb[k] = P.amb_sum[k] * mat->ambient[k];
                    if (COUNT) { cnt.rays += P.n_lights; cnt.shadow_rays += P.n_lights; cnt.pad += P.n_lights - P.n_lit_batch; }
                }
                if (P.n_lit_batch == 2) shade_small<2, COUNT>(P, w, h, mat, view, shade, dif, spc, cnt);
                else if (P.n_lit_batch == 3) shade_small<3, COUNT>(P, w, h, mat, view, shade, dif, spc, cnt);
                else if (P.n_lit_batch == 1) shade_small<1, COUNT>(P, w, h, mat, view, shade, dif, spc, cnt);
            } else {
            if (one_batch) {
                if (shade) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) amb[k] = P.amb_sum[k] * mat->ambient[k];
                    if (COUNT) { cnt.rays += P.n_lights; cnt.shadow_rays += P.n_lights; cnt.pad += P.n_lights - P.n_lit_batch; }
                }
            } else {
                // ambient term of every light, in light order (glsl:801)
                for (int l = 0; l < P.n_lights; ++l) {
                    if (shade) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) amb[k] = madd(P.lights[l].ambient[k], mat->ambient[k], amb[k]);
                        if (COUNT) { cnt.rays++; cnt.shadow_rays++; }
                    }
                }
            }
            // shadow rays toward the lit lights, kShadowBatch at a time (the light list is warp-uniform)
            int l = 0;
            while (one_batch ? l == 0 : l < P.n_lights) {
                int li[kShadowBatch];
                int n_slots = 0;
                if (one_batch) {
                    li[0] = P.lit[0]; li[1] = P.lit[1]; li[2] = P.lit[2];
                    n_slots = P.n_lit_batch;
                    l = P.n_lights;
                } else {
                    while (l < P.n_lights && n_slots < kShadowBatch) {
                        if (P.lights[l].position[3] != 0.0f) {
                            if (n_slots == 0) li[0] = l; else if (n_slots == 1) li[1] = l; else li[2] = l;
                            ++n_slots;
                        } else if (COUNT && shade) {
                            cnt.pad++;
                        }
                        ++l;
                    }
                }
                if (n_slots == 0) break;
                V3 so = mk(0.0f, 0.0f, 0.0f), sd[kShadowBatch];
                float sb[kShadowBatch];
                if (shade) so = axpy3(h.n, 0.01f, h.p);                                  // glsl:808
#pragma unroll
                for (int r = 0; r < kShadowBatch; ++r) {
                    if (shade && r < n_slots) {
                        const RtLightDev* __restrict__ L = P.lights + li[r];
                        sd[r] = mk(L->position[0] - h.p.x, L->position[1] - h.p.y, L->position[2] - h.p.z);  // glsl:809
                        sb[r] = 1.0f;                                               // glsl:816: shadowed iff closest t < 1
                    } else {
                        sd[r] = mk(0.0f, 0.0f, 1.0f);
                        sb[r] = 0.0f;
                    }
                }
                bool traced = false;
                if (REPACK && CULL && root_sb) {
#pragma unroll
                    for (int r = 0; r < kShadowBatch; ++r) sb[r] = sb[r] >= 1.0f ? first_sb[r] : sb[r];
                    traced = true;
                } else if (REPACK && CULL && P.packet) {
                    // coherent hit points (the usual case for the shadow rays of primary hits): packet tests
                    V3 lp[kShadowBatch];
#pragma unroll
                    for (int r = 0; r < kShadowBatch; ++r) {
                        const RtLightDev* __restrict__ L = P.lights + li[r < n_slots ? r : 0];
                        lp[r] = mk(L->position[0], L->position[1], L->position[2]);
                    }
                    Cone cone[kShadowBatch];
                    unsigned on;
                    // One pass when the warp's hit points are coherent; else (hit points of bounce rays scattered
                    // over the scene) groups of neighbouring hit points around a leader, one pass per group; lanes
                    // left over after kMaxGroups, or whose group sits too close to a light, take the per-ray loops.
                    const float rg2 = pk_pad2 * 2.5e5f * P.group_k;
                    bool todo = shade, left = false;
                    for (int round = 0; round < kMaxGroups; ++round) {
                        const unsigned rem = __ballot_sync(kFull, todo);
                        if (rem == 0u) break;
                        bool in_g = todo;
                        if (round > 0) {
                            const int lead = __ffs(rem) - 1;
                            const V3 dl = mk(h.p.x - __shfl_sync(kFull, h.p.x, lead), h.p.y - __shfl_sync(kFull, h.p.y, lead),
                                             h.p.z - __shfl_sync(kFull, h.p.z, lead));
                            in_g = todo && (w.lane == lead || dot3(dl, dl) <= rg2);
                        }
                        float sbg[kShadowBatch];
#pragma unroll
                        for (int r = 0; r < kShadowBatch; ++r) sbg[r] = in_g ? sb[r] : 0.0f;
                        if (shadow_cones(w, lp, sd, sbg, cone, on)) {
                            int unused[kShadowBatch] = {-1, -1, -1};
                            V3 so3[kShadowBatch] = {so, so, so};
                            packet_trace<SMEM, COUNT, kShadowBatch>(P, w, cone, on, 0.0101f, 1.0f, pk_pad1, pk_pad2, true, so3, sd, sbg, unused, cnt);
                            if (COUNT && w.lane == 0 && round > 0) cnt.shadow_group_passes++;      // round 0 is the whole warp
                            if (in_g) {
#pragma unroll
                                for (int r = 0; r < kShadowBatch; ++r) sb[r] = sbg[r];
                            }
                            todo = todo && !in_g;
                        } else if (round > 0) {
                            left = left || in_g;
                            todo = todo && !in_g;
                        }
                        if (!(P.group_k > 0.0f)) break;         // grouping switched off: the whole warp or nothing
                    }
                    // lanes already traced and not occluded are traced again by the per-ray loops below (same outcome)
                    traced = !__any_sync(kFull, todo || left);
                    if (COUNT && !traced && w.lane == 0) cnt.loop_passes++;
                }
                if (traced) {}
                else if (REPACK && CULL && P.n_lit > 0) trace_shadow<SMEM, REPACK, COUNT, CULL, true>(P, w, so, sd, sb, n_slots, cnt);
                else trace_shadow<SMEM, REPACK, COUNT, CULL, false>(P, w, so, sd, sb, n_slots, cnt);
#pragma unroll
                for (int r = 0; r < kShadowBatch; ++r) {
                    if (shade && r < n_slots && !(sb[r] < 1.0f)) {
                        if (COUNT) cnt.shaded_lights++;
                        phong_light(P.lights + li[r], mat, sd[r], h.n, view, dif, spc);
                    }
                }
            }
        }   // general path
        }
        if (hit) {
            if (shade) color = phong_accumulate(color, wp, amb, dif, spc, mat);                     // glsl:837-839
            // PHASES 2-3: each child ray is pushed with its product weight
            if (refracted) {                                                                        // glsl:1010-1027
                const V3 ro = axpy3(h.n, -0.001f, h.p);
                // glsl:1013-1020: ratio = 1 / ior, inverted again when leaving the object; both IEEE
                // quotients are per-object constants, built by the prologue
                const float ratio = h.inside ? mat->ior_back : mat->inv_ior;
                stack.push(ro, refract3(d[0], h.n, ratio), wgt * kt, depth - 1);
            }
            if (reflected) {                                                                        // glsl:988-994
                const V3 ro = axpy3(h.n, 0.001f, h.p);
                stack.push(ro, reflect3(d[0], h.n), wr, depth - 1);
            }
        }
        // Runaway guard (glsl:1076-1101): red iff the stack machine needs >= 10001 iterations.
        if (iters >= 10001) stack.clear();
    }
    if (iters >= 10001) { if (COUNT) cnt.red_pixels++; return mk(1.0f, 0.0f, 0.0f); }
    return color;
}

// -------------------------------------------------------------------------
// recursive_raytrace (glsl:1071-1105) for the THREE samples of a lane (one per 8x4 tile of the work item) in
// lockstep: large scenes with bounce rays.  trace_sample() runs the three bounce trees of a lane one after the
// other, so a grouping round sees 32 rays; here every iteration pops one pending ray per (lane, tile) -- up to 96 --
// and groups them across lanes AND tiles, for the closest-hit packets (secondary_packets' rule: direction within
// P.group_cos of a leader's, origin within sqrt(rg2) of it) and, light by light, for the shadow packets of the hit
// points (one cone from the light around a group of hit points, as the root phase does).  Every pass still scans
// all N objects; there are about half as many passes per ray.  Per-ray loops take what the groups leave over.
// Same nodes in the same per-pixel order, same exact tests (candidate_test is order-independent), same shading
// arithmetic (phong_light / phong_accumulate) as trace_sample: frames are byte-identical (no_cull bit 6).
// -------------------------------------------------------------------------
constexpr int kMaxGroups3 = 20;
template <bool SMEM, bool COUNT>
__device__ __forceinline__ void trace_sample3(const RtParams& P, WarpCtx& w, const bool (&valid)[kShadowBatch], V3 org,
                                              const V3 (&dir)[kShadowBatch], const float (&first_t)[kShadowBatch],
                                              const int (&first_idx)[kShadowBatch], bool have_sb,
                                              const float (&first_sb)[kShadowBatch][kShadowBatch], float pk_pad1, float pk_pad2,
                                              RtCounters& cnt, V3 (&color)[kShadowBatch]) {
    constexpr int Q = kShadowBatch;
    constexpr bool REPACK = true, CULL = true;
    RayStack<true> stack[Q];
    int iters[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) { stack[q].init(valid[q], org, dir[q], P.max_depth); color[q] = mk(0.0f, 0.0f, 0.0f); iters[q] = 0; }
    const float rg2 = pk_pad2 * 2.5e5f * P.group_k;
    const int n_lit = P.n_lit_batch;                      // 0..3 (checked by the caller)
    bool first_now = true;

    while (__any_sync(kFull, stack[0].pending() || stack[1].pending() || stack[2].pending())) {
        bool act[Q];
        V3 o[Q], d[Q];
        float wgt[Q], best[Q];
        int depth[Q], bidx[Q];
#pragma unroll
        for (int q = 0; q < Q; ++q) {
            act[q] = stack[q].pending();
            wgt[q] = 0.0f; depth[q] = 0; bidx[q] = -1;
            if (act[q]) { stack[q].pop(o[q], d[q], wgt[q], depth[q]); best[q] = kFar; }
            else { o[q] = mk(0.0f, 0.0f, 0.0f); d[q] = mk(0.0f, 0.0f, 1.0f); best[q] = 0.0f; }
        }
        // ---- PHASE 0 (glsl:956-967): closest hits
        if (first_now) {                                  // the root rays were traced by the caller
#pragma unroll
            for (int q = 0; q < Q; ++q) if (act[q]) { best[q] = first_t[q]; bidx[q] = first_idx[q]; }
        } else {
            bool todo[Q];
            V3 u[Q];
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                const float l2 = dot3(d[q], d[q]);
                todo[q] = act[q] && l2 > 0.0f && l2 < 1e30f;        // zero / NaN direction (refract()'s TIR): misses everything
                u[q] = d[q] * rsqrtf(l2);
            }
            for (int round = 0; round < kMaxGroups3; ++round) {
                const unsigned r0 = __ballot_sync(kFull, todo[0]), r1 = __ballot_sync(kFull, todo[1]), r2 = __ballot_sync(kFull, todo[2]);
                if ((r0 | r1 | r2) == 0u) break;
                const int ql = r0 ? 0 : (r1 ? 1 : 2);                  // the leader: first pending (tile, lane)
                const int lead = __ffs(r0 ? r0 : (r1 ? r1 : r2)) - 1;
                const V3 us = ql == 0 ? u[0] : (ql == 1 ? u[1] : u[2]);
                const V3 os = ql == 0 ? o[0] : (ql == 1 ? o[1] : o[2]);
                const V3 lu = mk(__shfl_sync(kFull, us.x, lead), __shfl_sync(kFull, us.y, lead), __shfl_sync(kFull, us.z, lead));
                const V3 lo = mk(__shfl_sync(kFull, os.x, lead), __shfl_sync(kFull, os.y, lead), __shfl_sync(kFull, os.z, lead));
                bool in_g[Q];
                float acc[7] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    const V3 dl = o[q] - lo;
                    in_g[q] = todo[q] && ((q == ql && w.lane == lead) || (dot3(u[q], lu) >= P.group_cos && dot3(dl, dl) <= rg2));
                    if (in_g[q]) { acc[0] += u[q].x; acc[1] += u[q].y; acc[2] += u[q].z; acc[3] += o[q].x; acc[4] += o[q].y; acc[5] += o[q].z; acc[6] += 1.0f; }
                }
#pragma unroll
                for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
                    for (int c = 0; c < 7; ++c) acc[c] += __shfl_xor_sync(kFull, acc[c], off);
                }
                Cone gc[1];
                const float inv_n = 1.0f / acc[6];
                gc[0].apex = mk(acc[3] * inv_n, acc[4] * inv_n, acc[5] * inv_n);
                gc[0].axis = normalize3(mk(acc[0], acc[1], acc[2]));
                float s2 = 0.0f, ro2 = 0.0f;
                unsigned want = 0u;
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    const V3 x = cross3(gc[0].axis, u[q]);
                    const V3 da = o[q] - gc[0].apex;
                    if (in_g[q]) {
                        s2 = fmaxf(s2, fminf(__fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z))), 1.0f));
                        ro2 = fmaxf(ro2, dot3(da, da));
                    }
                    if (__any_sync(kFull, in_g[q])) want |= 1u << q;
                }
                cone_close(gc[0], warp_max_nonneg(s2), 1e18f);
                const float grow = __fmaf_rn(sqrt_approx(warp_max_nonneg(ro2)), 1.001f, pk_pad1);
                float b3[Q]; int i3[Q];
#pragma unroll
                for (int q = 0; q < Q; ++q) { b3[q] = in_g[q] ? kFar : 0.0f; i3[q] = -1; }
                packet_trace<SMEM, COUNT, 1>(P, w, gc, 1u, grow, 1e-45f, pk_pad1, pk_pad2, false, o, d, b3, i3, cnt, want);
                if (COUNT && w.lane == 0) cnt.group_passes++;
#pragma unroll
                for (int q = 0; q < Q; ++q) if (in_g[q]) { best[q] = b3[q]; bidx[q] = i3[q]; todo[q] = false; }
            }
#pragma unroll 1
            for (int q = 0; q < Q; ++q) {                             // what the groups left over: per-ray loops
                const bool tq = q == 0 ? todo[0] : (q == 1 ? todo[1] : todo[2]);
                if (!__any_sync(kFull, tq)) continue;
                const V3 oq = q == 0 ? o[0] : (q == 1 ? o[1] : o[2]), dq = q == 0 ? d[0] : (q == 1 ? d[1] : d[2]);
                float b1 = tq ? kFar : 0.0f; int i1 = -1;
                if (COUNT && w.lane == 0) cnt.loop_passes++;
                trace_closest<SMEM, REPACK, COUNT, CULL>(P, w, oq, dq, b1, i1, cnt);
                if (tq) { if (q == 0) { best[0] = b1; bidx[0] = i1; } else if (q == 1) { best[1] = b1; bidx[1] = i1; } else { best[2] = b1; bidx[2] = i1; } }
            }
        }
        // ---- hit records, bounce weights (glsl:979-1054: final = mix(mix(phong, refl, kr), refr, kt) is linear)
        V3 hp[Q], hn[Q], so[Q], view[Q];
        float wp[Q], wr[Q], wt[Q], ratio[Q];
        int obj[Q];
        bool shade[Q], reflected[Q], refracted[Q], hit[Q];
#pragma unroll 1
        for (int q = 0; q < Q; ++q) {
            hit[q] = act[q] && bidx[q] >= 0;
            if (COUNT && act[q]) { cnt.rays++; cnt.nodes++; cnt.hit_nodes += hit[q]; }
            if (act[q]) iters[q] += hit[q] ? 6 : 1;
            hp[q] = hn[q] = so[q] = view[q] = mk(0.0f, 0.0f, 0.0f);
            wp[q] = wr[q] = wt[q] = ratio[q] = 0.0f; obj[q] = 0;
            shade[q] = reflected[q] = refracted[q] = false;
            if (hit[q]) {
                const Hit h = resolve_hit<SMEM, CULL>(P, w, o[q], d[q], best[q], bidx[q]);
                const RtCold* __restrict__ mat = P.cold + h.obj;
                const float kr = mat->reflectivity, kt = mat->transparency;
                reflected[q] = depth[q] > 0 && kr > 0.0f;
                refracted[q] = depth[q] > 0 && kt > 0.0f;
                float a = wgt[q], b = wgt[q] * kr;
                if (reflected[q]) a = a * (1.0f - kr);
                if (refracted[q]) { a = a * (1.0f - kt); b = b * (1.0f - kt); }
                wp[q] = a; wr[q] = b; wt[q] = wgt[q] * kt;
                ratio[q] = h.inside ? mat->ior_back : mat->inv_ior;
                hp[q] = h.p; hn[q] = h.n; obj[q] = h.obj;
                shade[q] = a != 0.0f;
                if (COUNT && !shade[q]) { cnt.rays += P.n_lights; cnt.shadow_rays += P.n_lights; cnt.pad += P.n_lights; }
                if (COUNT && shade[q]) { cnt.rays += P.n_lights; cnt.shadow_rays += P.n_lights; cnt.pad += P.n_lights - n_lit; }
                if (shade[q]) { view[q] = normalize3(-d[q]); so[q] = axpy3(h.n, 0.01f, h.p); }       // glsl:827, 808
            }
        }
        if (COUNT) cnt.warp_iters += Q;

        // ---- PHASE 1: shadow rays (glsl:807-819), light by light: sb[q][k] < 1 iff occluded
        float sb[Q][kShadowBatch];
        bool left[Q] = {false, false, false};
#pragma unroll
        for (int q = 0; q < Q; ++q)
#pragma unroll
            for (int k = 0; k < kShadowBatch; ++k) sb[q][k] = (shade[q] && k < n_lit) ? ((first_now && have_sb) ? first_sb[q][k] : 1.0f) : 0.0f;
        if (!(first_now && have_sb) && __any_sync(kFull, shade[0] || shade[1] || shade[2])) {
#pragma unroll 1
            for (int k = 0; k < n_lit; ++k) {
                const RtLightDev* __restrict__ L = P.lights + P.lit[k];
                const V3 lp = mk(L->position[0], L->position[1], L->position[2]);
                V3 sd3[Q], uu[Q];
                float ll[Q], sbk[Q];
                bool todo[Q];
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    sd3[q] = shade[q] ? mk(lp.x - hp[q].x, lp.y - hp[q].y, lp.z - hp[q].z) : mk(0.0f, 0.0f, 1.0f);      // glsl:809
                    ll[q] = __fmaf_rn(sd3[q].x, sd3[q].x, __fmaf_rn(sd3[q].y, sd3[q].y, __fmul_rn(sd3[q].z, sd3[q].z)));
                    uu[q] = sd3[q] * -rsqrtf(ll[q]);                  // from the light toward the hit point
                    todo[q] = shade[q];
                    sbk[q] = shade[q] ? 1.0f : 0.0f;
                }
                for (int round = 0; round < kMaxGroups3; ++round) {
                    const unsigned r0 = __ballot_sync(kFull, todo[0]), r1 = __ballot_sync(kFull, todo[1]), r2 = __ballot_sync(kFull, todo[2]);
                    if ((r0 | r1 | r2) == 0u) break;
                    bool in_g[Q] = {todo[0], todo[1], todo[2]};       // round 0: every hit point of the work item
                    if (round > 0) {                                  // then neighbours of a leader
                        const int ql = r0 ? 0 : (r1 ? 1 : 2);
                        const int lead = __ffs(r0 ? r0 : (r1 ? r1 : r2)) - 1;
                        const V3 ps = ql == 0 ? hp[0] : (ql == 1 ? hp[1] : hp[2]);
                        const V3 lh = mk(__shfl_sync(kFull, ps.x, lead), __shfl_sync(kFull, ps.y, lead), __shfl_sync(kFull, ps.z, lead));
#pragma unroll
                        for (int q = 0; q < Q; ++q) {
                            const V3 dl = hp[q] - lh;
                            in_g[q] = todo[q] && ((q == ql && w.lane == lead) || dot3(dl, dl) <= rg2);
                        }
                    }
                    V3 sm = mk(0.0f, 0.0f, 0.0f);
#pragma unroll
                    for (int q = 0; q < Q; ++q) if (in_g[q]) sm = sm + uu[q];
#pragma unroll
                    for (int off = 16; off > 0; off >>= 1) {
                        sm.x += __shfl_xor_sync(kFull, sm.x, off);
                        sm.y += __shfl_xor_sync(kFull, sm.y, off);
                        sm.z += __shfl_xor_sync(kFull, sm.z, off);
                    }
                    Cone lc[1];
                    lc[0].apex = lp;
                    lc[0].axis = normalize3(sm);
                    float s2 = 0.0f, lmax = 0.0f;
                    bool in_cone = true;
                    unsigned want = 0u;
#pragma unroll
                    for (int q = 0; q < Q; ++q) {
                        const V3 x = cross3(lc[0].axis, uu[q]);
                        if (in_g[q]) {
                            s2 = fmaxf(s2, fminf(__fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z))), 1.0f));
                            lmax = fmaxf(lmax, ll[q]);
                            in_cone = in_cone && dot3(lc[0].axis, uu[q]) >= kConeMinCos;            // NaN -> false
                        }
                        if (__any_sync(kFull, in_g[q])) want |= 1u << q;
                    }
                    if (__all_sync(kFull, in_cone)) {
                        cone_close(lc[0], warp_max_nonneg(s2), __fmaf_rn(sqrt_approx(warp_max_nonneg(lmax)), 1.0001f, 0.0101f));
                        float sbg[Q]; int unused[Q] = {-1, -1, -1};
#pragma unroll
                        for (int q = 0; q < Q; ++q) sbg[q] = in_g[q] ? 1.0f : 0.0f;
                        packet_trace<SMEM, COUNT, 1>(P, w, lc, 1u, 0.0101f, 1.0f, pk_pad1, pk_pad2, false, so, sd3, sbg, unused, cnt, want);
                        if (COUNT && w.lane == 0 && round > 0) cnt.shadow_group_passes++;
#pragma unroll
                        for (int q = 0; q < Q; ++q) if (in_g[q]) { sbk[q] = sbg[q]; todo[q] = false; }
                    } else if (round > 0) {                           // this neighbourhood is too wide seen from the light
#pragma unroll
                        for (int q = 0; q < Q; ++q) if (in_g[q]) { left[q] = true; todo[q] = false; }
                    }
                }
#pragma unroll
                for (int q = 0; q < Q; ++q) {
                    left[q] = left[q] || todo[q];
                    if (k == 0) sb[q][0] = sbk[q]; else if (k == 1) sb[q][1] = sbk[q]; else sb[q][2] = sbk[q];
                }
            }
#pragma unroll 1
            for (int q = 0; q < Q; ++q) {                             // leftovers: the per-ray loops, all lights of the hit point
                const bool lq = q == 0 ? left[0] : (q == 1 ? left[1] : left[2]);
                if (!__any_sync(kFull, lq)) continue;
                if (COUNT && w.lane == 0) cnt.loop_passes++;
                V3 sdq[kShadowBatch]; float sbq[kShadowBatch];
#pragma unroll
                for (int k = 0; k < kShadowBatch; ++k) {
                    const RtLightDev* __restrict__ L = P.lights + P.lit[k < n_lit ? k : 0];
                    const bool on = shade[q] && k < n_lit;
                    sdq[k] = on ? mk(L->position[0] - hp[q].x, L->position[1] - hp[q].y, L->position[2] - hp[q].z) : mk(0.0f, 0.0f, 1.0f);
                    sbq[k] = on ? 1.0f : 0.0f;
                }
                trace_shadow<SMEM, REPACK, COUNT, CULL, false>(P, w, so[q], sdq, sbq, n_lit, cnt);
#pragma unroll
                for (int k = 0; k < kShadowBatch; ++k) if (shade[q] && k < n_lit) sb[q][k] = sbq[k];      // same outcome for rays traced before
            }
        }

        // ---- shading (glsl:789-840) and PHASES 2-3: each child ray is pushed with its product weight
#pragma unroll 1
        for (int q = 0; q < Q; ++q) {
            if (hit[q]) {
                const RtCold* __restrict__ mat = P.cold + obj[q];
                if (shade[q]) {
                    float amb[4], dif[4] = {0, 0, 0, 0}, spc[4] = {0, 0, 0, 0};
#pragma unroll
                    for (int k = 0; k < 4; ++k) amb[k] = P.amb_sum[k] * mat->ambient[k];
                    for (int k = 0; k < n_lit; ++k) {
                        if (!(sb[q][k] < 1.0f)) {
                            if (COUNT) cnt.shaded_lights++;
                            const RtLightDev* __restrict__ L = P.lights + P.lit[k];
                            phong_light(L, mat, mk(L->position[0] - hp[q].x, L->position[1] - hp[q].y, L->position[2] - hp[q].z), hn[q], view[q], dif, spc);
                        }
                    }
                    color[q] = phong_accumulate(color[q], wp[q], amb, dif, spc, mat);
                }
                if (refracted[q]) stack[q].push(axpy3(hn[q], -0.001f, hp[q]), refract3(d[q], hn[q], ratio[q]), wt[q], depth[q] - 1);   // glsl:1010-1027
                if (reflected[q]) stack[q].push(axpy3(hn[q], 0.001f, hp[q]), reflect3(d[q], hn[q]), wr[q], depth[q] - 1);              // glsl:988-994
            }
            if (iters[q] >= 10001) stack[q].clear();                  // runaway guard (glsl:1076-1101)
        }
        first_now = false;
    }
#pragma unroll
    for (int q = 0; q < Q; ++q) if (iters[q] >= 10001) { if (COUNT) cnt.red_pixels++; color[q] = mk(1.0f, 0.0f, 0.0f); }
}

// -------------------------------------------------------------------------
// The kernel.  grid = persistent CTAs; each warp walks 8x4-pixel tiles of all
// frames (tile index -> frame, tile row, tile column).
// -------------------------------------------------------------------------
template <bool SMEM, bool REPACK, bool COUNT, bool CULL, bool BOUNCE = true>
__global__ void __launch_bounds__(kThreads, REPACK ? (CULL ? RT_MIN_BLOCKS_CULL : RT_MIN_BLOCKS) : RT_MIN_BLOCKS_SMALL) render_kernel(const __grid_constant__ RtParams P) {
    // dynamic shared memory: [hot scene (resident only)] [per-warp scratch] [per-warp tile buffers + barriers (streaming only)]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    WarpCtx w;
    w.lane = lane;
    w.phase = 0u;
    size_t off4 = 0;                         // in float4 units
    if (SMEM) {
        // stage the hot scene once per CTA (persistent: once per SM-slot per launch)
        const int nb4 = P.n_box * box_tile_f4<CULL>(), ns4 = P.n_sph;
        const float4* __restrict__ box_src = CULL ? P.box_cull : P.box_hot;
        for (int i = threadIdx.x; i < nb4; i += kThreads) smem_f4[i] = __ldg(box_src + i);
        for (int i = threadIdx.x; i < ns4; i += kThreads) smem_f4[nb4 + i] = __ldg(P.sph_hot + i);
        w.box_off4 = 0u;
        w.sph_off4 = (uint32_t)nb4;
        off4 = (size_t)nb4 + ns4;
        w.lrec_off4 = (uint32_t)off4;
        if (CULL && P.n_lit > 0) {
            const int nl4 = n_lrec_tiles(P) * P.n_lit * kTileBox;
            for (int i = threadIdx.x; i < nl4; i += kThreads) smem_f4[off4 + i] = __ldg(P.obj_lrec + i);
            off4 += (size_t)nl4;
        }
    } else {
        w.box_off4 = w.sph_off4 = w.lrec_off4 = 0u;
    }
    {
        float4* scratch = smem_f4 + off4 + (size_t)warp * (kScratchBytes / 16);
        w.queue = reinterpret_cast<float*>(scratch);
        w.occ = reinterpret_cast<int*>(scratch + kQueueRays * 2);
        off4 += (size_t)kWarps * (kScratchBytes / 16);
    }
    w.buf_off4[0] = w.buf_off4[1] = 0u;
    w.buf_addr[0] = w.buf_addr[1] = w.bar_addr[0] = w.bar_addr[1] = 0u;
    if (!SMEM) {
        w.buf_off4[0] = (uint32_t)(off4 + (size_t)warp * (2 * kTileBytes / 16));
        w.buf_off4[1] = w.buf_off4[0] + kTileBytes / 16;
        off4 += (size_t)kWarps * (2 * kTileBytes / 16);
        uint64_t* bars = reinterpret_cast<uint64_t*>(smem_f4 + off4) + warp * 2;
        w.buf_addr[0] = smem_addr(smem_f4 + w.buf_off4[0]); w.buf_addr[1] = smem_addr(smem_f4 + w.buf_off4[1]);
        w.bar_addr[0] = smem_addr(bars); w.bar_addr[1] = smem_addr(bars + 1);
        if (lane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(w.bar_addr[0]) : "memory");
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(w.bar_addr[1]) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    __syncthreads();

    const int warp_global = blockIdx.x * kWarps + warp;
    const int n_warps = gridDim.x * kWarps;
    const int tiles_per_frame = P.tiles_x * P.tiles_y;
    const long long n_tiles = (long long)tiles_per_frame * P.n_frames;
    const int lx = lane & 7, ly = lane >> 3;

    // packet tests: radius pads that scale with the scene extent E (see rt_prologue_math.h)
    float pk_pad1 = 0.0f, pk_pad2 = 0.0f;
    if (CULL && P.packet) {
        const float E = __ldg(P.extent);
        pk_pad1 = 2e-6f * E;
        pk_pad2 = 4e-6f * E * E;
    }

    RtCounters cnt;
    cnt.packet_tests = cnt.group_passes = cnt.loop_passes = cnt.candidates = cnt.shadow_group_passes = 0ull;
    cnt.rays = cnt.nodes = cnt.hit_nodes = cnt.shadow_rays = cnt.box_tests = cnt.sphere_tests = cnt.red_pixels = cnt.pad = 0ull;
    cnt.cull_tests = cnt.warp_iters = cnt.shaded_lights = 0ull;

    // Work distribution: the first round of tiles is static (tile = global warp id); after that
    // each warp pulls the next tile index from a global counter (zeroed by the host before the
    // launch), so long and short tiles (glass vs wall, occluded vs open) even out.
    // Large scenes: a work item is NP = 3 horizontally adjacent 8x4 tiles, one pixel of each per lane.
    // The three primary rays of a lane start at the camera (glsl:391), so they go through the object
    // loop together with the origin moved into each primitive's frame once (like the shadow rays of
    // one hit point); P.tiles_x counts work items.
    constexpr int NP = REPACK ? kShadowBatch : 1;
    for (long long tile = warp_global; tile < n_tiles;) {
        const int frame = (int)(tile / tiles_per_frame);
        const int tin = (int)(tile - (long long)frame * tiles_per_frame);
        const int ty = tin / P.tiles_x, tx = tin - ty * P.tiles_x;
        const int py = P.y0 + ty * 4 + ly;
        int px[NP]; bool valid[NP];
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            px[q] = (tx * NP + q) * 8 + lx;
            valid[q] = px[q] < P.width && py < P.y1;
        }
        const RtCameraDev* __restrict__ cam = P.cams ? P.cams + frame : &P.cam0;
        const V3 org = mk(cam->position[0], cam->position[1], cam->position[2]);   // glsl:391

        V3 sum[NP];
#pragma unroll
        for (int q = 0; q < NP; ++q) sum[q] = mk(0.0f, 0.0f, 0.0f);
        for (int s = 0; s < P.spp; ++s) {
            V3 dir[kShadowBatch];
#pragma unroll
            for (int q = 0; q < NP; ++q) {
                float jx, jy;
                jitter(P, px[q], py, s, jx, jy);
                dir[q] = primary_dir(P, cam, px[q], py, jx, jy);
            }
            const float no_sb[kShadowBatch] = {0.0f, 0.0f, 0.0f};
            if (NP == 1) {
                sum[0] = sum[0] + trace_sample<SMEM, REPACK, COUNT, CULL, BOUNCE>(P, w, valid[0], org, dir[0], false, 0.0f, -1, false, no_sb, 0.0f, 0.0f, cnt);
            } else {
                // PHASE 0 of the three root frames (glsl:956-967), shared origin
                V3 o3[kShadowBatch];
                float best[kShadowBatch]; int bidx[kShadowBatch];
#pragma unroll
                for (int q = 0; q < kShadowBatch; ++q) {
                    o3[q] = org; bidx[q] = -1;
                    best[q] = valid[q] ? kFar : 0.0f;
                }
                // The work item's primary rays as one packet: a cone from the camera around the direction
                // through the middle of the 24x4-pixel strip.
                Cone pc[1];
                bool packet = false;
                if (CULL && P.packet) {
                    pc[0].apex = org;
                    pc[0].axis = primary_dir(P, cam, tx * (8 * NP) + 4 * NP, P.y0 + ty * 4 + 2, 0.0f, 0.0f);
                    float s2 = 0.0f;
                    bool in_cone = true;
#pragma unroll
                    for (int q = 0; q < kShadowBatch; ++q) {
                        const V3 x = cross3(pc[0].axis, dir[q]);
                        const float s2q = __fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z)));
                        if (valid[q]) {
                            s2 = fmaxf(s2, fminf(s2q, 1.0f));
                            in_cone = in_cone && dot3(pc[0].axis, dir[q]) >= kConeMinCos;
                        }
                    }
                    packet = __all_sync(kFull, in_cone);
                    cone_close(pc[0], warp_max_nonneg(s2), kFar * 1.001f);
                }
                if (CULL && packet) {
                    packet_trace<SMEM, COUNT, 1>(P, w, pc, 1u, 0.0f, 1e-45f, pk_pad1, pk_pad2, true, o3, dir, best, bidx, cnt);
                } else {
                    const int n_obj_tiles = n_box_tiles<CULL>(P) + n_sph_tiles(P);
                    tile_begin<SMEM, CULL>(P, w);
                    for (int t = 0; t < n_obj_tiles; ++t) {
                        const float4* recs = tile_acquire<SMEM, CULL>(P, w, t, n_obj_tiles);
                        any_tile<kShadowBatch, false, true, COUNT, CULL>(P, recs, t, o3, dir, best, bidx, cnt);
                    }
                }
                // The shadow rays of the three root hits, light by light: pass k traces, for every lane, the rays
                // from its three hit points toward lit light k (glsl:807-819) as ONE packet -- a cone from the
                // light around the work item's 96 hit points -- instead of three packets of three cones each.
                // so / sd are built exactly as the shading code builds them (resolve_hit is deterministic).
                bool have_sb = false;
                float rsb[kShadowBatch][kShadowBatch];       // [tile q][light slot]
#pragma unroll
                for (int q = 0; q < kShadowBatch; ++q)
#pragma unroll
                    for (int k = 0; k < kShadowBatch; ++k) rsb[q][k] = 0.0f;
                if (CULL && packet && !RT_STRICT && P.n_lit_batch > 0) {
                    V3 hp[kShadowBatch], so3[kShadowBatch];
                    bool sh[kShadowBatch];
#pragma unroll 1
                    for (int q = 0; q < kShadowBatch; ++q) {
                        const V3 dq = q == 0 ? dir[0] : (q == 1 ? dir[1] : dir[2]);
                        const float tq = q == 0 ? best[0] : (q == 1 ? best[1] : best[2]);
                        const int iq = q == 0 ? bidx[0] : (q == 1 ? bidx[1] : bidx[2]);
                        const bool vq = q == 0 ? valid[0] : (q == 1 ? valid[1] : valid[2]);
                        bool shade = false;
                        V3 p = mk(0.0f, 0.0f, 0.0f), so = mk(0.0f, 0.0f, 0.0f);
                        if (vq && iq >= 0) {
                            const Hit h = resolve_hit<SMEM, CULL>(P, w, org, dq, tq, iq);
                            const RtCold* __restrict__ mat = P.cold + h.obj;
                            // the weight of the Phong term, as trace_sample computes it for the root node
                            float wp = 1.0f;
                            if (P.max_depth > 0 && mat->reflectivity > 0.0f) wp = wp * (1.0f - mat->reflectivity);
                            if (P.max_depth > 0 && mat->transparency > 0.0f) wp = wp * (1.0f - mat->transparency);
                            shade = wp != 0.0f;
                            p = h.p;
                            so = axpy3(h.n, 0.01f, h.p);                                            // glsl:808
                        }
                        if (q == 0) { hp[0] = p; so3[0] = so; sh[0] = shade; }
                        else if (q == 1) { hp[1] = p; so3[1] = so; sh[1] = shade; }
                        else { hp[2] = p; so3[2] = so; sh[2] = shade; }
                    }
                    have_sb = true;
                    if (__any_sync(kFull, sh[0] || sh[1] || sh[2])) {
#pragma unroll 1
                        for (int k = 0; k < P.n_lit_batch; ++k) {
                            const RtLightDev* __restrict__ L = P.lights + P.lit[k];
                            const V3 lp = mk(L->position[0], L->position[1], L->position[2]);
                            V3 sd3[kShadowBatch], u[kShadowBatch];
                            float l2[kShadowBatch], sbk[kShadowBatch];
                            V3 sum = mk(0.0f, 0.0f, 0.0f);
#pragma unroll
                            for (int q = 0; q < kShadowBatch; ++q) {
                                sd3[q] = sh[q] ? mk(lp.x - hp[q].x, lp.y - hp[q].y, lp.z - hp[q].z) : mk(0.0f, 0.0f, 1.0f);   // glsl:809
                                sbk[q] = sh[q] ? 1.0f : 0.0f;
                                l2[q] = __fmaf_rn(sd3[q].x, sd3[q].x, __fmaf_rn(sd3[q].y, sd3[q].y, __fmul_rn(sd3[q].z, sd3[q].z)));
                                u[q] = sd3[q] * -rsqrtf(l2[q]);                                // from the light toward the hit point
                                if (sh[q]) sum = sum + u[q];
                            }
#pragma unroll
                            for (int off = 16; off > 0; off >>= 1) {
                                sum.x += __shfl_xor_sync(kFull, sum.x, off);
                                sum.y += __shfl_xor_sync(kFull, sum.y, off);
                                sum.z += __shfl_xor_sync(kFull, sum.z, off);
                            }
                            Cone lc[1];
                            lc[0].apex = lp;
                            lc[0].axis = normalize3(sum);
                            float s2 = 0.0f, lmax = 0.0f;
                            bool in_cone = true;
#pragma unroll
                            for (int q = 0; q < kShadowBatch; ++q) {
                                const V3 x = cross3(lc[0].axis, u[q]);
                                const float s2q = __fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z)));
                                if (sh[q]) {
                                    s2 = fmaxf(s2, fminf(s2q, 1.0f));
                                    lmax = fmaxf(lmax, l2[q]);
                                    in_cone = in_cone && dot3(lc[0].axis, u[q]) >= kConeMinCos;     // NaN -> false
                                }
                            }
                            if (!__all_sync(kFull, in_cone)) { have_sb = false; break; }          // wide packet: per tile below
                            cone_close(lc[0], warp_max_nonneg(s2), __fmaf_rn(sqrt_approx(warp_max_nonneg(lmax)), 1.0001f, 0.0101f));
                            int unused[kShadowBatch] = {-1, -1, -1};
                            packet_trace<SMEM, COUNT, 1>(P, w, lc, 1u, 0.0101f, 1.0f, pk_pad1, pk_pad2, false, so3, sd3, sbk, unused, cnt);
#pragma unroll
                            for (int q = 0; q < kShadowBatch; ++q) {
                                if (k == 0) rsb[q][0] = sbk[q]; else if (k == 1) rsb[q][1] = sbk[q]; else rsb[q][2] = sbk[q];
                            }
                        }
                    }
                }
                // bounce rays of a large scene: the three trees of a lane in lockstep, rays grouped over all 96 (trace_sample3)
                bool lockstep = false;
                if constexpr (BOUNCE && CULL && REPACK && !RT_STRICT) {
                    lockstep = P.packet && P.group_k > 0.0f && P.n_lit_batch >= 0;       // warp-uniform
                    if (lockstep) {
                        const bool v3[kShadowBatch] = {valid[0], valid[NP > 1 ? 1 : 0], valid[NP > 2 ? 2 : 0]};
                        V3 c3[kShadowBatch];
                        trace_sample3<SMEM, COUNT>(P, w, v3, org, dir, best, bidx, have_sb, rsb, pk_pad1, pk_pad2, cnt, c3);
#pragma unroll
                        for (int q = 0; q < NP; ++q) sum[q] = sum[q] + c3[q];
                    }
                }
                // one copy of the bounce machinery in the instruction stream: rolled loop, selected operands
#pragma unroll 1
                for (int q = 0; q < (lockstep ? 0 : NP); ++q) {
                    const V3 dq = q == 0 ? dir[0] : (q == 1 ? dir[1] : dir[2]);
                    const float tq = q == 0 ? best[0] : (q == 1 ? best[1] : best[2]);
                    const int iq = q == 0 ? bidx[0] : (q == 1 ? bidx[1] : bidx[2]);
                    const bool vq = q == 0 ? valid[0] : (q == 1 ? valid[1] : valid[2]);
                    float fsb[kShadowBatch];
#pragma unroll
                    for (int k = 0; k < kShadowBatch; ++k) fsb[k] = q == 0 ? rsb[0][k] : (q == 1 ? rsb[1][k] : rsb[2][k]);
                    const V3 c = trace_sample<SMEM, REPACK, COUNT, CULL, BOUNCE>(P, w, vq, org, dq, true, tq, iq, have_sb, fsb, pk_pad1, pk_pad2, cnt);
                    if (q == 0) sum[0] = sum[0] + c;
                    else if (q == 1) sum[1] = sum[1] + c;
                    else sum[2] = sum[2] + c;
                }
            }
        }
        uint8_t* frame_out = P.out + (long long)frame * P.frame_stride;
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const float fs = (float)P.spp;
            const V3 fc = mk(fdiv(sum[q].x, fs), fdiv(sum[q].y, fs), fdiv(sum[q].z, fs));
            // vec4(final_color, 0.0) -> RGBA8 (glsl:404): R | G<<8 | B<<16 | A(=0)<<24
            const uint32_t word = unorm8(fc.x) | (unorm8(fc.y) << 8) | (unorm8(fc.z) << 16);
            const long long pix = (long long)(py - P.out_row0) * P.width + px[q];
            // 16-byte stores: lanes 4k..4k+3 hold 4 horizontally adjacent pixels
            const uint32_t w1 = __shfl_down_sync(kFull, word, 1);
            const uint32_t w2 = __shfl_down_sync(kFull, word, 2);
            const uint32_t w3 = __shfl_down_sync(kFull, word, 3);
            const bool vec_ok = ((P.width & 3) == 0) && ((tx * NP + q) * 8 + 8 <= P.width) &&
                                ((reinterpret_cast<uintptr_t>(frame_out) & 15) == 0);
            if (vec_ok) {
                if ((lane & 3) == 0 && py < P.y1)
                    *reinterpret_cast<uint4*>(frame_out + pix * 4) = make_uint4(word, w1, w2, w3);
            } else if (valid[q]) {
                *reinterpret_cast<uint32_t*>(frame_out + pix * 4) = word;
            }
        }
        unsigned long long next = 0ull;
        if (lane == 0) next = atomicAdd(P.tile_counter, 1ull);
        tile = (long long)n_warps + (long long)__shfl_sync(kFull, next, 0);
    }

    if (COUNT) {
        // warp-reduce then one atomic per warp per counter
        if (lane != 0) cnt.warp_iters = 0ull;    // a per-warp count, kept by lane 0
        unsigned long long v[16] = {cnt.rays, cnt.nodes, cnt.hit_nodes, cnt.shadow_rays, cnt.box_tests, cnt.sphere_tests,
                                    cnt.red_pixels, cnt.pad, cnt.cull_tests, cnt.warp_iters, cnt.packet_tests, cnt.group_passes, cnt.loop_passes,
                                    cnt.candidates, cnt.shadow_group_passes, cnt.shaded_lights};
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            unsigned long long x = v[k];
            for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(kFull, x, off);
            if (lane == 0 && x) atomicAdd(reinterpret_cast<unsigned long long*>(P.counters) + k, x);
        }
    }
}

#if !RT_STRICT
#include "rt_kernel_block.cuh"
#endif

}  // namespace RT_NS

// Launchers with C linkage-free, namespace-qualified names; called from rt_host.cpp.
namespace RT_NS {
inline size_t hot_smem_bytes(const RtParams& P, bool cull) {
    size_t f4 = (size_t)P.n_box * (cull ? 1 : RT_BOX_F4) + (size_t)P.n_sph;
    if (cull && P.n_lit > 0) f4 += (size_t)((P.n_box + P.n_sph + kTileBox - 1) / kTileBox) * P.n_lit * kTileBox;
    return f4 * sizeof(float4);
}
inline size_t scratch_smem_bytes() { return (size_t)kWarps * kScratchBytes; }
inline size_t stream_smem_bytes() { return (size_t)kWarps * (2 * kTileBytes + 16); }
size_t resident_limit_bytes(int max_smem_optin, bool cull);
cudaError_t launch(const RtParams& P, int sm_count, bool count, cudaStream_t stream);
cudaError_t configure(int device, int* max_smem_optin);
}
