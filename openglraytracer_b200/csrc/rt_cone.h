/*
 * rt_cone.h -- the packet tests of the large-scene kernels: cone, reach, fan plane, ball (DESIGN.md 4, items 15, 19,
 * 20).  Included by rt_kernel.cuh inside the kernel namespace.  The same text compiles as plain C++ for
 * tests/cone_host.cpp, which checks on the CPU that the tests are conservative (a packet test may only reject an
 * object that none of the packet's rays can hit): there __fmaf_rn is fmaf, __fmul_rn a product (built with
 * -ffp-contract=off) and the approximate sqrt / rsqrt are the IEEE ones (1-2 ulp apart, far inside the pads).
 */
#pragma once
#ifdef __CUDACC__
#define RTC_FN __device__ __forceinline__
#else
#include <cmath>
#define RTC_FN inline
struct V3 { float x, y, z; };
struct float4 { float x, y, z, w; };
inline V3 mk(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
inline V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline float __fmaf_rn(float a, float b, float c) { return std::fmaf(a, b, c); }
inline float __fmul_rn(float a, float b) { return a * b; }
inline float sqrt_approx(float x) { return std::sqrt(x); }
inline float rsqrtf(float x) { return 1.0f / std::sqrt(x); }
#endif

struct Cone {
    V3 apex, axis;    // cone_open(): axis = 0, c2 = 0: no angular bound (everything within `reach` of the apex passes the cone part)
    float s, c2;      // padded sin(theta), 1 - s^2
    float reach;      // no ray of the packet extends further than this from the apex
    // Fan: the packet's rays also stay within `hw` of the plane through the apex with unit normal `nrm`.  Shadow
    // rays of hit points strung out along a thin view frustum subtend a wide angle at the light but lie
    // close to the plane through the light and the frustum's axis.  hw = 3e38: no plane.
    V3 nrm;
    float hw;
};

RTC_FN V3 cross3(V3 a, V3 b) {
    return mk(__fmaf_rn(a.y, b.z, -__fmul_rn(a.z, b.y)), __fmaf_rn(a.z, b.x, -__fmul_rn(a.x, b.z)),
              __fmaf_rn(a.x, b.y, -__fmul_rn(a.y, b.x)));
}
// c = (centre, .), rp = padded bounding radius (NaN for a box that can never be hit -> false)
template <bool SLAB = false>
RTC_FN bool cone_pass(const Cone& k, float4 c, float rp) {
    const V3 v = mk(c.x - k.apex.x, c.y - k.apex.y, c.z - k.apex.z);
    const float h = __fmaf_rn(v.x, k.axis.x, __fmaf_rn(v.y, k.axis.y, __fmul_rn(v.z, k.axis.z)));
    const float v2 = __fmaf_rn(v.x, v.x, __fmaf_rn(v.y, v.y, __fmul_rn(v.z, v.z)));
    // squared distance to the axis as |v|^2 - h^2 (7 instructions fewer than |v x axis|^2; this test is what the
    // large-scene kernels spend their time in).  The difference cancels, so it is lowered by 2e-6 |v|^2 -- above the
    // rounding of v2, of h (|axis| = 1 +- 4e-7) and of the FMA together (1.4e-6 |v|^2): never larger than the true
    // value, so the test stays conservative; far from the apex it is a little less tight (0.0014 |v| in p).
    const float p2 = __fmaf_rn(-2e-6f, v2, __fmaf_rn(-h, h, v2));
    const float rhs = __fmaf_rn(h, k.s, rp);
    const float far = k.reach + rp;
    bool pass = rhs >= 0.0f && __fmul_rn(p2, k.c2) <= __fmul_rn(rhs, rhs) && v2 <= __fmul_rn(far, far);
    if (SLAB) {
        const float sl = fabsf(__fmaf_rn(v.x, k.nrm.x, __fmaf_rn(v.y, k.nrm.y, __fmul_rn(v.z, k.nrm.z))));
        pass = pass && sl <= k.hw + rp;
    }
    return pass;
}
// Finishes a cone from the warp's maxima: s2 = max |axis x u|^2 over the packet's unit directions u.
RTC_FN void cone_close(Cone& k, float s2max, float reach) {
    k.s = __fmaf_rn(sqrt_approx(s2max), 1.0005f, 4e-6f);
    k.c2 = fmaxf(1.0f - k.s * k.s, 0.0f);
    k.reach = reach;
    k.nrm = mk(0.0f, 0.0f, 0.0f);
    k.hw = 3e38f;
}
// No angular bound: everything within `reach` of the apex (and within the fan plane's slab) passes.
RTC_FN void cone_open(Cone& k) {
    k.axis = mk(0.0f, 0.0f, 0.0f);
    k.s = 1.0f;
    k.c2 = 0.0f;
}
// Unit normal of the plane through a light and a view axis (a = the axis, b = eye - light); zero when the light
// sits on the axis (then the cone is narrow anyway).
RTC_FN V3 fan_normal(V3 a, V3 b) {
    const V3 x = cross3(a, b);
    const float l2 = __fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z)));
    return l2 > 1e-12f ? x * rsqrtf(l2) : mk(0.0f, 0.0f, 0.0f);
}
// hmax = max |nrm . (p - apex)| over the packet's hit points; pad1 = 2e-6 E covers the rounding of both dot products
RTC_FN void cone_fan(Cone& k, V3 nrm, float hmax, float pad1) {
    const bool ok = (nrm.x != 0.0f || nrm.y != 0.0f || nrm.z != 0.0f) && hmax < 1e30f;      // NaN -> no plane
    k.nrm = ok ? nrm : mk(0.0f, 0.0f, 0.0f);
    k.hw = ok ? __fmaf_rn(hmax, 1.0005f, 4.0f * pad1) : 3e38f;
}
// Padded bounding radius of a record.  Boxes: w = padded squared circumradius (prologue).  Spheres: w = r*r,
// padded here by the cancellation noise of the shader's quadratic (pad2 = 4e-6 E^2) and the coordinate
// rounding (pad1 = 2e-6 E).  `grow`: what the packet's rays may be off their cone by (shadow rays: 0.0101).
RTC_FN float packet_radius(float w, bool sph, float grow, float pad1, float pad2) {
    float rp = sqrt_approx(sph ? w + pad2 : w);
    if (sph) rp = __fmaf_rn(rp, 1.0005f, pad1);
    return __fmaf_rn(rp, 1.000001f, grow);
}
// Groups of objects (rt_groups.cu): g = (centre m, R) with R >= |c - m| + packet_radius(member, grow = 0) for every
// member (c, .).  group_radius is the radius to put into cone_pass / ball_pass in a member's place: whenever one of
// those tests passes for a member with the same `grow`, it passes for (m, group_radius).  Why: each test compares a
// function of the centre that changes by at most |c - m| between c and m (the distance to the cone's surface
// p cos - h sin, -h sin, |v|, |v . nrm|: gradients of length <= 1) with the radius, so R + grow on the right side
// covers the move.  Two things are not 1-Lipschitz and are covered by the 0.0015 |v| term: the lowering of p^2 by
// 2e-6 |v|^2 in cone_pass, which may let a member pass whose true distance to the cone is up to 0.0014 |v_member|
// (<= 0.0014 (|v| + R)) beyond its radius, and the rounding of h, v2 (a few 1e-7 |v|).  R itself carries a relative
// 1e-4 and an absolute 1e-5 E from the builder.  tests/cone_host.cpp checks the implication on random packets.
RTC_FN float group_radius(V3 apex, float4 g, float grow) {
    const V3 v = mk(g.x - apex.x, g.y - apex.y, g.z - apex.z);
    const float v2 = __fmaf_rn(v.x, v.x, __fmaf_rn(v.y, v.y, __fmul_rn(v.z, v.z)));
    return __fmaf_rn(sqrt_approx(v2), 0.0015f, __fmaf_rn(g.w, 1.0016f, grow));
}
// Rays that start within `radius` of the apex and run toward it stay inside that ball: the bounding sphere
// (c, rp) can only matter if it reaches into the ball.
RTC_FN bool ball_pass(V3 apex, float radius, float4 c, float rp) {
    const V3 v = mk(c.x - apex.x, c.y - apex.y, c.z - apex.z);
    const float far = radius + rp;
    return __fmaf_rn(v.x, v.x, __fmaf_rn(v.y, v.y, __fmul_rn(v.z, v.z))) <= __fmul_rn(far, far);
}
