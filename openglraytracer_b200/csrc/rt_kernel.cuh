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
// Same association as the oracle's pinned dot(): ((x*x + y*y) + z*z)
__device__ __forceinline__ float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 normalize3(V3 v) { float l = sqrtf(dot3(v, v)); return mk(v.x / l, v.y / l, v.z / l); }
__device__ __forceinline__ V3 reflect3(V3 I, V3 N) { float d2 = 2.0f * dot3(N, I); return I - d2 * N; }
__device__ __forceinline__ V3 refract3(V3 I, V3 N, float eta) {
    float ni = dot3(N, I);
    float k = 1.0f - eta * eta * (1.0f - ni * ni);
    if (k < 0.0f) return mk(0.0f, 0.0f, 0.0f);
    return eta * I - (eta * ni + sqrtf(k)) * N;
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
// Ray / sphere test (glsl:585-625).  c = (centre, r*r).  Returns t (or <= 0 / NaN
// for "no usable hit") and whether the origin is inside.
// -------------------------------------------------------------------------
struct RayQ { float qa, iqa; };   // per-ray invariants of the quadratic (fast path)

#if RT_STRICT
__device__ __forceinline__ float sphere_t(float4 c, V3 o, V3 d, RayQ, bool& inside) {
    V3 oc = mk(o.x - c.x, o.y - c.y, o.z - c.z);
    float qa = dot3(d, d);
    float qb = dot3(2.0f * d, oc);
    float qc = dot3(oc, oc) - c.w;
    float qd = qb * qb - 4.0f * qa * qc;
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
__device__ __forceinline__ float sphere_t(float4 c, V3 o, V3 d, RayQ q, bool& inside) {
    float ocx = o.x - c.x, ocy = o.y - c.y, ocz = o.z - c.z;
    float hb = __fmaf_rn(d.x, ocx, __fmaf_rn(d.y, ocy, __fmul_rn(d.z, ocz)));       // qb / 2
    float qc = __fmaf_rn(ocx, ocx, __fmaf_rn(ocy, ocy, __fmaf_rn(ocz, ocz, -c.w)));
    float disc = __fmaf_rn(hb, hb, -__fmul_rn(q.qa, qc));                           // qd / 4
    float sq = sqrt_approx(fmaxf(disc, 0.0f));
    float tn = __fmul_rn(-hb - sq, q.iqa);
    float tf = __fmul_rn(sq - hb, q.iqa);
    inside = tn < 0.0f;
    float t = inside ? tf : tn;
    return (disc < 0.0f) ? -1.0f : t;      // t_far < 0 gives t < 0: rejected by the caller's t > 0
}
#endif

__device__ __forceinline__ RayQ ray_q(V3 d) {
    RayQ q;
    q.qa = __fmaf_rn(d.x, d.x, __fmaf_rn(d.y, d.y, __fmul_rn(d.z, d.z)));
    q.iqa = 1.0f / q.qa;
    return q;
}

// -------------------------------------------------------------------------
// get_closest_collision (glsl:738-782) for R rays of this lane at once.
// best[r] comes in as t_max (10000 for closest-hit rays, 1 for shadow rays,
// 0 for an inactive slot) and goes out as the smallest accepted t; bidx[r] is
// the hot-list position of the winner (boxes 0..n_box-1, spheres n_box..) or -1.
// The loop trip count is warp-uniform.
//   SHARED : all R rays start at o[0] (the shadow rays of one hit point,
//            glsl:808) -> the origin is moved into the box frame once per box.
//   ANYHIT : shadow semantics (glsl:816: in shadow iff closest t < 1, i.e. iff
//            ANY object is hit at 0 < t < 1): no index, warp-vote early exit.
// Most tests miss, so the accept logic sits behind one warp vote per box.
// -------------------------------------------------------------------------
template <int R, bool SMEM, bool ANYHIT, bool SHARED, bool COUNT>
__device__ __forceinline__ void trace(const RtParams& P, const float4* __restrict__ box_hot,
                                      const float4* __restrict__ sph_hot,
                                      const V3 (&o)[R], const V3 (&d)[R], float (&best)[R], int (&bidx)[R],
                                      unsigned long long& n_box_tests, unsigned long long& n_sph_tests) {
    const int n_box = P.n_box, n_sph = P.n_sph;
    bool stop = false;
#pragma unroll 2
    for (int i = 0; i < n_box; ++i) {
        float4 rec[RT_BOX_F4];
#pragma unroll
        for (int k = 0; k < RT_BOX_F4; ++k) rec[k] = SMEM ? box_hot[i * RT_BOX_F4 + k] : __ldg(box_hot + i * RT_BOX_F4 + k);
        float tn[R], tf[R];
        bool cand = false;
        V3 so;
        if (SHARED) so = box_origin(rec, o[0]);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (COUNT) n_box_tests += ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f);
            if (!SHARED) so = box_origin(rec, o[r]);
            Slabs s; V3 ld;
            box_slabs(rec, so, d[r], tn[r], tf[r], s, ld);
            cand = cand || !box_reject(tn[r], tf[r]);
        }
        if (__any_sync(kFull, cand)) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float t = box_pick(tn[r], tf[r]);
                if (!box_reject(tn[r], tf[r]) && t > 0.0f && t < best[r]) { best[r] = t; if (!ANYHIT) bidx[r] = i; }
            }
        }
        if (ANYHIT && (i & 15) == 15) {
            bool done = true;
#pragma unroll
            for (int r = 0; r < R; ++r) done = done && (best[r] < 1.0f);
            if (__all_sync(kFull, done)) { stop = true; break; }
        }
    }
    if (ANYHIT && stop) return;

    RayQ q[R];
#pragma unroll
    for (int r = 0; r < R; ++r) q[r] = ray_q(d[r]);
#pragma unroll 2
    for (int j = 0; j < n_sph; ++j) {
        const float4 c = SMEM ? sph_hot[j] : __ldg(sph_hot + j);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (COUNT) n_sph_tests += ANYHIT ? (best[r] >= 1.0f) : (best[r] > 0.0f);
            bool inside;
            float t = sphere_t(c, o[SHARED ? 0 : r], d[r], q[r], inside);
            if (t > 0.0f && t < best[r]) { best[r] = t; if (!ANYHIT) bidx[r] = n_box + j; }
            else if (!ANYHIT && t == best[r] && bidx[r] >= 0 && bidx[r] < n_box) {
                // exact tie with a box: the lower ORIGINAL index wins (glsl:773 strict '<' in index order)
                if (__ldg(P.sph_obj + j) < __ldg(P.box_obj + bidx[r])) bidx[r] = n_box + j;
            }
        }
        if (ANYHIT && (j & 15) == 15) {
            bool done = true;
#pragma unroll
            for (int r = 0; r < R; ++r) done = done && (best[r] < 1.0f);
            if (__all_sync(kFull, done)) break;
        }
    }
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

// Re-evaluates the winning primitive to produce p, n, inside (glsl:627-637 sphere,
// glsl:686-721 box).  t is the value the trace loop accepted.
template <bool SMEM>
__device__ __forceinline__ Hit resolve_hit(const RtParams& P, const float4* __restrict__ box_hot,
                                           const float4* __restrict__ sph_hot, V3 o, V3 d, float t, int bidx) {
    Hit h;
    if (bidx >= P.n_box) {
        const int j = bidx - P.n_box;
        const float4 c = SMEM ? sph_hot[j] : __ldg(sph_hot + j);
        bool inside;
        (void)sphere_t(c, o, d, ray_q(d), inside);
        h.obj = __ldg(P.sph_obj + j);
        h.inside = inside;
        h.p = o + t * d;                                            // glsl:628
        h.n = normalize3(mk(h.p.x - c.x, h.p.y - c.y, h.p.z - c.z));  // glsl:631
        if (inside) h.n = h.n * -1.0f;                              // glsl:634-637
        return h;
    }
    h.obj = __ldg(P.box_obj + bidx);
    const RtCold* __restrict__ c = P.cold + h.obj;
    float4 rec[RT_BOX_F4];
#pragma unroll
    for (int k = 0; k < RT_BOX_F4; ++k) rec[k] = SMEM ? box_hot[bidx * RT_BOX_F4 + k] : __ldg(box_hot + bidx * RT_BOX_F4 + k);
    float tn, tf; Slabs s; V3 ld;
    box_slabs(rec, box_origin(rec, o), d, tn, tf, s, ld);
    h.inside = tn < 0.0f;                                           // glsl:690-696
    const float by = h.inside ? s.t2y : s.t1y, bz = h.inside ? s.t2z : s.t1z;
    int face = 0;                                                   // glsl:699-704
    if (t == by) face = 1;
    else if (t == bz) face = 2;
    // ray in the box's (unscaled) local space, as the shader computes it (glsl:655-656)
    const float* w = c->w2l;
    V3 rs, rd;
    rs.x = ((w[0] * o.x + w[1] * o.y) + w[2] * o.z) + w[3] * 1.0f;
    rs.y = ((w[4] * o.x + w[5] * o.y) + w[6] * o.z) + w[7] * 1.0f;
    rs.z = ((w[8] * o.x + w[9] * o.y) + w[10] * o.z) + w[11] * 1.0f;
    rd.x = ((w[0] * d.x + w[1] * d.y) + w[2] * d.z) + w[3] * 0.0f;
    rd.y = ((w[4] * d.x + w[5] * d.y) + w[6] * d.z) + w[7] * 0.0f;
    rd.z = ((w[8] * d.x + w[9] * d.y) + w[10] * d.z) + w[11] * 0.0f;
    const float rdf = face == 0 ? rd.x : (face == 1 ? rd.y : rd.z);
    V3 nl = mk(face == 0 ? 1.0f : 0.0f, face == 1 ? 1.0f : 0.0f, face == 2 ? 1.0f : 0.0f);   // glsl:707-708
    if (rdf > 0.0f) nl = nl * -1.0f;                                // glsl:712-715
    const float* m = c->nmat;                                       // glsl:718, column-major
    h.n.x = (m[0] * nl.x + m[3] * nl.y) + m[6] * nl.z;
    h.n.y = (m[1] * nl.x + m[4] * nl.y) + m[7] * nl.z;
    h.n.z = (m[2] * nl.x + m[5] * nl.y) + m[8] * nl.z;
    const V3 ql = rs + t * rd;                                      // glsl:721
    const float* l = c->l2w;
    h.p.x = ((l[0] * ql.x + l[1] * ql.y) + l[2] * ql.z) + l[3] * 1.0f;
    h.p.y = ((l[4] * ql.x + l[5] * ql.y) + l[6] * ql.z) + l[7] * 1.0f;
    h.p.z = ((l[8] * ql.x + l[9] * ql.y) + l[10] * ql.z) + l[11] * 1.0f;
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
        s[r] = ((m[r] * vx + m[4 + r] * vy) + m[8 + r] * 0.5f) + m[12 + r] * 1.0f;
        e[r] = ((m[r] * vx + m[4 + r] * vy) + m[8 + r] * 1.0f) + m[12 + r] * 1.0f;
    }
    V3 ws = mk(s[0] / s[3], s[1] / s[3], s[2] / s[3]);
    V3 we = mk(e[0] / e[3], e[1] / e[3], e[2] / e[3]);
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
template <bool SMEM, bool COUNT>
__device__ __forceinline__ V3 trace_sample(const RtParams& P, const float4* __restrict__ box_hot,
                                           const float4* __restrict__ sph_hot, bool valid, V3 org, V3 dir,
                                           RtCounters& cnt) {
    // pending rays: origin, direction, weight, remaining depth
    float stk[kStack][8];
    int sp = 0;
    if (valid) {
        stk[0][0] = org.x; stk[0][1] = org.y; stk[0][2] = org.z;
        stk[0][3] = dir.x; stk[0][4] = dir.y; stk[0][5] = dir.z;
        stk[0][6] = 1.0f; stk[0][7] = __int_as_float(P.max_depth);
        sp = 1;
    }
    V3 color = mk(0.0f, 0.0f, 0.0f);
    int iters = 0;   // loop iterations the shader's stack machine would have spent (glsl:1081-1094)

    while (__any_sync(kFull, sp > 0)) {
        const bool act = sp > 0;
        V3 o[1], d[1];
        float w = 0.0f; int depth = 0;
        float best[1]; int bidx[1];
        bidx[0] = -1;
        if (act) {
            --sp;
            o[0] = mk(stk[sp][0], stk[sp][1], stk[sp][2]);
            d[0] = mk(stk[sp][3], stk[sp][4], stk[sp][5]);
            w = stk[sp][6]; depth = __float_as_int(stk[sp][7]);
            best[0] = kFar;
        } else {
            o[0] = mk(0.0f, 0.0f, 0.0f); d[0] = mk(0.0f, 0.0f, 1.0f);
            best[0] = 0.0f;
        }
        // PHASE 0 (glsl:956-967)
        trace<1, SMEM, false, false, COUNT>(P, box_hot, sph_hot, o, d, best, bidx, cnt.box_tests, cnt.sphere_tests);
        const bool hit = act && bidx[0] >= 0;
        if (COUNT && act) { cnt.rays++; cnt.nodes++; cnt.hit_nodes += hit; }
        if (act) iters += hit ? 6 : 1;

        Hit h;
        const RtCold* __restrict__ mat = P.cold;
        if (hit) {
            h = resolve_hit<SMEM>(P, box_hot, sph_hot, o[0], d[0], best[0], bidx[0]);
            mat = P.cold + h.obj;
        }
        // PHASE 1: ads_phong_lighting (glsl:789-840), shadow rays kShadowBatch at a time
        float amb[4] = {0, 0, 0, 0}, dif[4] = {0, 0, 0, 0}, spc[4] = {0, 0, 0, 0};
        V3 view = mk(0.0f, 0.0f, 0.0f);
        if (hit) view = normalize3(-d[0]);                          // glsl:827 normalize(-r.dir)
        if (__any_sync(kFull, hit)) {
            for (int l0 = 0; l0 < P.n_lights; l0 += kShadowBatch) {
                V3 so[kShadowBatch], sd[kShadowBatch];
                float sb[kShadowBatch]; int si[kShadowBatch];
#pragma unroll
                for (int r = 0; r < kShadowBatch; ++r) {
                    const int l = l0 + r;
                    si[r] = -1;
                    if (hit && l < P.n_lights) {
                        const RtLightDev* __restrict__ L = P.lights + l;
                        so[r] = h.p + h.n * 0.01f;                                  // glsl:808
                        sd[r] = mk(L->position[0] - h.p.x, L->position[1] - h.p.y, L->position[2] - h.p.z);  // glsl:809
                        sb[r] = 1.0f;                                               // glsl:816: shadowed iff closest t < 1
                    } else {
                        so[r] = mk(0.0f, 0.0f, 0.0f); sd[r] = mk(0.0f, 0.0f, 1.0f);
                        sb[r] = 0.0f;
                    }
                }
                trace<kShadowBatch, SMEM, true, true, COUNT>(P, box_hot, sph_hot, so, sd, sb, si, cnt.box_tests, cnt.sphere_tests);
#pragma unroll
                for (int r = 0; r < kShadowBatch; ++r) {
                    const int l = l0 + r;
                    if (hit && l < P.n_lights) {
                        if (COUNT) { cnt.rays++; cnt.shadow_rays++; }
                        const RtLightDev* __restrict__ L = P.lights + l;
#pragma unroll
                        for (int k = 0; k < 4; ++k) amb[k] += L->ambient[k] * mat->ambient[k];     // glsl:801
                        const bool in_shadow = sb[r] < 1.0f;
                        if (!in_shadow) {
                            const V3 light_dir = normalize3(sd[r]);                                  // glsl:804
                            const V3 light_ref = normalize3(reflect3(-light_dir, h.n));              // glsl:825
                            const float cos_theta = dot3(light_dir, h.n);                            // glsl:826
                            const float cos_phi = dot3(view, light_ref);                             // glsl:827
                            const float kd = fmaxf(cos_theta, 0.0f);
                            const float ks = powf(fmaxf(cos_phi, 0.0f), mat->shininess);             // glsl:832
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                dif[k] += L->diffuse[k] * mat->diffuse[k] * kd;                      // glsl:830
                                spc[k] += L->specular[k] * mat->specular[k] * ks;                    // glsl:832
                            }
                        }
                    }
                }
            }
        }
        if (hit) {
            float pc[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) pc[k] = ((amb[k] + dif[k]) + spc[k]) + mat->emissive[k];    // glsl:837
            const V3 phong = mk(pc[0] * pc[3], pc[1] * pc[3], pc[2] * pc[3]);                        // glsl:839
            // PHASES 2-4 (glsl:979-1054): final = mix(mix(phong, refl, kr), refr, kt) is linear in
            // its three colours, so each child is pushed with its product weight.
            const float kr = mat->reflectivity, kt = mat->transparency;
            const bool reflected = depth > 0 && kr > 0.0f;
            const bool refracted = depth > 0 && kt > 0.0f;
            float wp = w, wr = w * kr;
            if (reflected) wp = wp * (1.0f - kr);
            if (refracted) { wp = wp * (1.0f - kt); wr = wr * (1.0f - kt); }
            color = color + wp * phong;
            if (refracted) {                                                                        // glsl:1010-1027
                const V3 ro = h.p - h.n * 0.001f;
                float ratio = 1.0f / mat->refraction_index;
                if (h.inside) ratio = 1.0f / ratio;
                const V3 rdv = refract3(d[0], h.n, ratio);
                stk[sp][0] = ro.x; stk[sp][1] = ro.y; stk[sp][2] = ro.z;
                stk[sp][3] = rdv.x; stk[sp][4] = rdv.y; stk[sp][5] = rdv.z;
                stk[sp][6] = w * kt; stk[sp][7] = __int_as_float(depth - 1);
                ++sp;
            }
            if (reflected) {                                                                        // glsl:988-994
                const V3 ro = h.p + h.n * 0.001f;
                const V3 rdv = reflect3(d[0], h.n);
                stk[sp][0] = ro.x; stk[sp][1] = ro.y; stk[sp][2] = ro.z;
                stk[sp][3] = rdv.x; stk[sp][4] = rdv.y; stk[sp][5] = rdv.z;
                stk[sp][6] = wr; stk[sp][7] = __int_as_float(depth - 1);
                ++sp;
            }
        }
        // Runaway guard (glsl:1076-1101): red iff the stack machine needs >= 10001 iterations.
        if (iters >= 10001) sp = 0;
    }
    if (iters >= 10001) { if (COUNT) cnt.red_pixels++; return mk(1.0f, 0.0f, 0.0f); }
    return color;
}

// -------------------------------------------------------------------------
// The kernel.  grid = persistent CTAs; each warp walks 8x4-pixel tiles of all
// frames (tile index -> frame, tile row, tile column).
// -------------------------------------------------------------------------
template <bool SMEM, bool COUNT>
__global__ void __launch_bounds__(kThreads, 2) render_kernel(const __grid_constant__ RtParams P) {
    extern __shared__ float4 smem_f4[];
    const float4* box_hot = P.box_hot;
    const float4* sph_hot = P.sph_hot;
    if (SMEM) {
        // stage the hot scene once per CTA (persistent: once per SM-slot per launch)
        const int nb4 = P.n_box * RT_BOX_F4, ns4 = P.n_sph;
        for (int i = threadIdx.x; i < nb4; i += kThreads) smem_f4[i] = __ldg(P.box_hot + i);
        for (int i = threadIdx.x; i < ns4; i += kThreads) smem_f4[nb4 + i] = __ldg(P.sph_hot + i);
        __syncthreads();
        box_hot = smem_f4;
        sph_hot = smem_f4 + nb4;
    }

    const int lane = threadIdx.x & 31;
    const int warp_global = blockIdx.x * kWarps + (threadIdx.x >> 5);
    const int n_warps = gridDim.x * kWarps;
    const int tiles_per_frame = P.tiles_x * P.tiles_y;
    const long long n_tiles = (long long)tiles_per_frame * P.n_frames;
    const int lx = lane & 7, ly = lane >> 3;

    RtCounters cnt;
    cnt.rays = cnt.nodes = cnt.hit_nodes = cnt.shadow_rays = cnt.box_tests = cnt.sphere_tests = cnt.red_pixels = 0ull;

    for (long long tile = warp_global; tile < n_tiles; tile += n_warps) {
        const int frame = (int)(tile / tiles_per_frame);
        const int tin = (int)(tile - (long long)frame * tiles_per_frame);
        const int ty = tin / P.tiles_x, tx = tin - ty * P.tiles_x;
        const int px = tx * 8 + lx;
        const int py = P.y0 + ty * 4 + ly;
        const bool valid = px < P.width && py < P.y1;
        const RtCameraDev* __restrict__ cam = P.cams + frame;
        const V3 org = mk(cam->position[0], cam->position[1], cam->position[2]);   // glsl:391

        V3 sum = mk(0.0f, 0.0f, 0.0f);
        for (int s = 0; s < P.spp; ++s) {
            float jx, jy;
            jitter(P, px, py, s, jx, jy);
            const V3 dir = primary_dir(P, cam, px, py, jx, jy);
            const V3 c = trace_sample<SMEM, COUNT>(P, box_hot, sph_hot, valid, org, dir, cnt);
            sum = sum + c;
        }
        const float fs = (float)P.spp;
        const V3 fc = mk(sum.x / fs, sum.y / fs, sum.z / fs);
        // vec4(final_color, 0.0) -> RGBA8 (glsl:404): R | G<<8 | B<<16 | A(=0)<<24
        const uint32_t word = unorm8(fc.x) | (unorm8(fc.y) << 8) | (unorm8(fc.z) << 16);

        uint8_t* frame_out = P.out + (long long)frame * P.frame_stride;
        const long long pix = (long long)(py - P.out_row0) * P.width + px;
        // 16-byte stores: lanes 4k..4k+3 hold 4 horizontally adjacent pixels
        const uint32_t w1 = __shfl_down_sync(kFull, word, 1);
        const uint32_t w2 = __shfl_down_sync(kFull, word, 2);
        const uint32_t w3 = __shfl_down_sync(kFull, word, 3);
        const bool vec_ok = ((P.width & 3) == 0) && (tx * 8 + 8 <= P.width) &&
                            ((reinterpret_cast<uintptr_t>(frame_out) & 15) == 0);
        if (vec_ok) {
            if ((lane & 3) == 0 && py < P.y1)
                *reinterpret_cast<uint4*>(frame_out + pix * 4) = make_uint4(word, w1, w2, w3);
        } else if (valid) {
            *reinterpret_cast<uint32_t*>(frame_out + pix * 4) = word;
        }
    }

    if (COUNT) {
        // warp-reduce then one atomic per warp per counter
        unsigned long long v[7] = {cnt.rays, cnt.nodes, cnt.hit_nodes, cnt.shadow_rays, cnt.box_tests, cnt.sphere_tests, cnt.red_pixels};
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            unsigned long long x = v[k];
            for (int off = 16; off > 0; off >>= 1) x += __shfl_down_sync(kFull, x, off);
            if (lane == 0 && x) atomicAdd(reinterpret_cast<unsigned long long*>(P.counters) + k, x);
        }
    }
}

}  // namespace RT_NS

// Launchers with C linkage-free, namespace-qualified names; called from rt_host.cpp.
namespace RT_NS {
inline size_t hot_smem_bytes(const RtParams& P) { return ((size_t)P.n_box * RT_BOX_F4 + (size_t)P.n_sph) * sizeof(float4); }
cudaError_t launch(const RtParams& P, int grid, bool count, cudaStream_t stream);
cudaError_t configure(int device, int* max_smem_optin);
}
