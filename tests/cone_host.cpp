// CPU property test of the packet tests in openglraytracer_b200/csrc/rt_cone.h (the text the kernels compile):
// a packet test may only reject an object that NONE of the packet's rays can hit.  Random packets are built the
// way the kernels build them (rt_kernel.cuh: packet cones around a mean direction, rt_kernel_block.cuh: light
// cones with fan planes, secondary_packets: groups with scattered origins) and checked against an exact
// double-precision segment / sphere test.  Prints "<cases> <objects that must pass> <violations>".
// Built by tests/test_cone_host.py: g++ -O2 -ffp-contract=off.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "rt_cone.h"

static uint64_t rng_state = 0x5CE7E001ull;
static double uni() {   // splitmix64 -> [0, 1)
    uint64_t z = (rng_state += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (double)(z >> 11) * (1.0 / 9007199254740992.0);
}
static double uni(double a, double b) { return a + (b - a) * uni(); }
struct D3 { double x, y, z; };
static D3 operator-(D3 a, D3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static D3 operator+(D3 a, D3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static D3 operator*(D3 a, double s) { return {a.x * s, a.y * s, a.z * s}; }
static double dot(D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static D3 unit(D3 a) { return a * (1.0 / std::sqrt(dot(a, a))); }
static D3 rnd_unit() {
    for (;;) {
        D3 v = {uni(-1, 1), uni(-1, 1), uni(-1, 1)};
        const double l = dot(v, v);
        if (l > 1e-3 && l <= 1.0) return unit(v);
    }
}
static V3 f3(D3 a) { return mk((float)a.x, (float)a.y, (float)a.z); }
static D3 d3(V3 a) { return {a.x, a.y, a.z}; }
// distance of c to the segment o + t d, t in [0, tmax]
static double seg_dist(D3 o, D3 d, double tmax, D3 c) {
    const double dd = dot(d, d);
    double t = dd > 0 ? dot(c - o, d) / dd : 0.0;
    t = t < 0 ? 0 : (t > tmax ? tmax : t);
    const D3 q = o + d * t - c;
    return std::sqrt(dot(q, q));
}
static float cross_s2(V3 a, V3 u) {
    const V3 x = cross3(a, u);
    return std::fmin(__fmaf_rn(x.x, x.x, __fmaf_rn(x.y, x.y, __fmul_rn(x.z, x.z))), 1.0f);
}
static V3 unit_sum(V3 s) { return s * rsqrtf(s.x * s.x + s.y * s.y + s.z * s.z); }

struct Packet {
    std::vector<D3> o, d;     // ray i: o + t d, t in [0, 1] (d carries the length)
    Cone k;
    float grow = 0.0f;
    bool slab = false;
};

// primary-like packet: one apex, directions within a random half-angle up to ~50 degrees, reach 10000 (glsl:740)
static void primary_packet(Packet& P, int n) {
    const D3 apex = {uni(-10, 10), uni(-10, 10), uni(-10, 10)};
    const D3 ax = rnd_unit();
    const double spread = uni(0.001, 0.85);
    V3 sum = mk(0, 0, 0);
    std::vector<V3> u(n);
    P.k.apex = f3(apex);
    for (int i = 0; i < n; ++i) {
        const D3 dir = unit(ax + rnd_unit() * (spread * uni()));
        u[i] = f3(dir);
        P.o[i] = d3(P.k.apex);            // the kernel's rays start at the FP32 apex
        P.d[i] = d3(u[i]) * 40.0;         // longer than the room
        sum = mk(sum.x + u[i].x, sum.y + u[i].y, sum.z + u[i].z);
    }
    P.k.axis = unit_sum(sum);
    float s2 = 0.0f;
    for (int i = 0; i < n; ++i) s2 = std::fmax(s2, cross_s2(P.k.axis, u[i]));
    cone_close(P.k, s2, 10000.0f * 1.001f);
}

// shadow packet of a block: hit points strung out along a view frustum, rays p + 0.01 n -> L (glsl:808-809), cone
// from the light (opened when wide), reach, fan plane through the light and the view axis (rt_kernel_block.cuh)
static void shadow_packet(Packet& P, int n, float pad1) {
    const V3 L = f3({uni(-9, 9), uni(-9, 9), uni(-9, 9)});
    const D3 eye = {uni(-10, 10), uni(-10, 10), uni(-10, 10)};
    const D3 view = rnd_unit();
    const double half = uni(0.002, 0.03);
    std::vector<V3> u(n), sd(n);
    V3 sum = mk(0, 0, 0);
    float lmax = 0.0f;
    for (int i = 0; i < n; ++i) {
        const D3 dir = unit(view + rnd_unit() * (half * uni()));
        const V3 p = f3(eye + dir * uni(0.5, 25.0));
        const V3 nr = f3(rnd_unit());
        sd[i] = mk(L.x - p.x, L.y - p.y, L.z - p.z);                                                    // glsl:809
        P.o[i] = d3(mk(__fmaf_rn(nr.x, 0.01f, p.x), __fmaf_rn(nr.y, 0.01f, p.y), __fmaf_rn(nr.z, 0.01f, p.z)));   // glsl:808
        P.d[i] = d3(sd[i]);
        const float l2 = __fmaf_rn(sd[i].x, sd[i].x, __fmaf_rn(sd[i].y, sd[i].y, __fmul_rn(sd[i].z, sd[i].z)));
        u[i] = sd[i] * -rsqrtf(l2);
        sum = mk(sum.x + u[i].x, sum.y + u[i].y, sum.z + u[i].z);
        lmax = std::fmax(lmax, l2);
    }
    P.k.apex = L;
    P.k.axis = unit_sum(sum);
    const V3 nrm = fan_normal(f3(view), mk((float)eye.x - L.x, (float)eye.y - L.y, (float)eye.z - L.z));
    float s2 = 0.0f, hmax = 0.0f;
    bool narrow = true;
    for (int i = 0; i < n; ++i) {
        s2 = std::fmax(s2, cross_s2(P.k.axis, u[i]));
        narrow = narrow && (P.k.axis.x * u[i].x + P.k.axis.y * u[i].y + P.k.axis.z * u[i].z) >= 0.1f;
        hmax = std::fmax(hmax, std::fabs(nrm.x * sd[i].x + nrm.y * sd[i].y + nrm.z * sd[i].z));
    }
    cone_close(P.k, s2, __fmaf_rn(sqrt_approx(lmax), 1.0001f, 0.0101f));
    if (!narrow) cone_open(P.k);
    cone_fan(P.k, nrm, hmax, pad1);
    P.grow = 0.0101f;
    P.slab = true;
}

// group of bounce rays: scattered origins, directions within 8 degrees, apex = mean origin, radii grown by the
// origin spread (rt_kernel.cuh: secondary_packets)
static void group_packet(Packet& P, int n, float pad1) {
    const D3 base = {uni(-9, 9), uni(-9, 9), uni(-9, 9)};
    const D3 ax = rnd_unit();
    const double rg = uni(0.01, 1.9);
    V3 sum = mk(0, 0, 0), so = mk(0, 0, 0);
    std::vector<V3> u(n), of(n);
    for (int i = 0; i < n; ++i) {
        u[i] = f3(unit(ax + rnd_unit() * (0.13 * uni())));
        of[i] = f3(base + rnd_unit() * (rg * uni()));
        P.o[i] = d3(of[i]);
        P.d[i] = d3(u[i]) * 40.0;
        sum = mk(sum.x + u[i].x, sum.y + u[i].y, sum.z + u[i].z);
        so = mk(so.x + of[i].x, so.y + of[i].y, so.z + of[i].z);
    }
    P.k.apex = so * (1.0f / (float)n);
    P.k.axis = unit_sum(sum);
    float s2 = 0.0f, ro2 = 0.0f;
    for (int i = 0; i < n; ++i) {
        s2 = std::fmax(s2, cross_s2(P.k.axis, u[i]));
        const V3 da = mk(of[i].x - P.k.apex.x, of[i].y - P.k.apex.y, of[i].z - P.k.apex.z);
        ro2 = std::fmax(ro2, da.x * da.x + da.y * da.y + da.z * da.z);
    }
    cone_close(P.k, s2, 1e18f);
    P.grow = __fmaf_rn(sqrt_approx(ro2), 1.001f, pad1);
}

int main(int argc, char** argv) {
    const int cases = argc > 1 ? atoi(argv[1]) : 2000;
    const double E = 19.05;                       // extent of the +-11 room
    const float pad1 = (float)(2e-6 * E), pad2 = (float)(4e-6 * E * E);
    long must = 0, bad = 0, member_pass = 0, group_bad = 0;
    for (int cs = 0; cs < cases; ++cs) {
        const int n_rays = 4 + (int)(uni() * 93);
        Packet P;
        P.o.resize(n_rays); P.d.resize(n_rays);
        if (cs % 3 == 0) primary_packet(P, n_rays);
        else if (cs % 3 == 1) shadow_packet(P, n_rays, pad1);
        else group_packet(P, n_rays, pad1);
        // objects: a mix of sizes; half of them placed next to a ray so that near misses and grazing hits are common
        std::vector<float4> recs;                  // every object's record and un-grown packet radius, for the groups below
        std::vector<float> rp0s;
        for (int j = 0; j < 400; ++j) {
            const double R = j % 4 == 0 ? uni(0.5, 3.0) : uni(0.02, 0.6);
            D3 c = {uni(-11, 11), uni(-11, 11), uni(-11, 11)};
            if (j % 2) {
                const int i = (int)(uni() * n_rays);
                c = P.o[i] + P.d[i] * uni() + rnd_unit() * (R * uni(0.5, 1.5));
            }
            const bool sph = j % 5 == 4;
            // the record the prologue writes: squared radius, padded for boxes (rt_prologue_math.h: padded_radius2)
            const double rp_box = R * 1.0005 + 2e-6 * E;
            const float4 rec = {(float)c.x, (float)c.y, (float)c.z, sph ? (float)(R * R) : (float)(rp_box * rp_box)};
            recs.push_back(rec);
            rp0s.push_back(packet_radius(rec.w, sph, 0.0f, pad1, pad2) / 1.000001f * 1.0000002f);   // its value before the last step, rounded up
            const D3 cf = {rec.x, rec.y, rec.z};
            bool hit = false;
            for (int i = 0; i < n_rays && !hit; ++i) hit = seg_dist(P.o[i], P.d[i], 1.0, cf) <= R;
            if (!hit) continue;
            ++must;
            const float rp = packet_radius(rec.w, sph, P.grow, pad1, pad2);
            const bool pass = P.slab ? cone_pass<true>(P.k, rec, rp) : cone_pass<false>(P.k, rec, rp);
            if (!pass) ++bad;
        }
        // groups (rt_groups.cu): objects of one cell of a 4 x 4 x 4 grid, 8 at a time; the sphere is built as the
        // kernel builds it (middle of the centres' bounding box, |c - m| and the sum rounded up by 1e-4, + 1e-5 E).
        // Whatever packet test passes for a member must pass for its group with group_radius.
        {
            std::vector<int> order(recs.size());
            for (size_t j = 0; j < order.size(); ++j) order[j] = (int)j;
            auto cell = [&](int j) {
                auto q = [](float v) { int c = (int)((v + 11.0f) * (4.0f / 22.0f)); return c < 0 ? 0 : (c > 3 ? 3 : c); };
                return (q(recs[j].x) * 4 + q(recs[j].y)) * 4 + q(recs[j].z);
            };
            std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return cell(a) < cell(b); });
            const float ball_r = (float)uni(0.05, 4.0);
            for (size_t g0 = 0; g0 < order.size(); g0 += 8) {
                const size_t g1 = std::min(order.size(), g0 + 8);
                float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
                for (size_t e = g0; e < g1; ++e) {
                    const float4 c = recs[order[e]];
                    lo[0] = std::fmin(lo[0], c.x); lo[1] = std::fmin(lo[1], c.y); lo[2] = std::fmin(lo[2], c.z);
                    hi[0] = std::fmax(hi[0], c.x); hi[1] = std::fmax(hi[1], c.y); hi[2] = std::fmax(hi[2], c.z);
                }
                const float mx = 0.5f * lo[0] + 0.5f * hi[0], my = 0.5f * lo[1] + 0.5f * hi[1], mz = 0.5f * lo[2] + 0.5f * hi[2];
                float r = 0.0f;
                for (size_t e = g0; e < g1; ++e) {
                    const float4 c = recs[order[e]];
                    const float dx = c.x - mx, dy = c.y - my, dz = c.z - mz;
                    r = std::fmax(r, __fmaf_rn(std::sqrt(dx * dx + dy * dy + dz * dz), 1.0001f, rp0s[order[e]]));
                }
                const float4 grp = {mx, my, mz, __fmaf_rn(r, 1.0001f, (float)(1e-5 * E))};
                const float rg = group_radius(P.k.apex, grp, P.grow);
                const bool g_cone = P.slab ? cone_pass<true>(P.k, grp, rg) : cone_pass<false>(P.k, grp, rg);
                const bool g_ball = ball_pass(P.k.apex, ball_r, grp, rg);
                for (size_t e = g0; e < g1; ++e) {
                    const int j = order[e];
                    const float rm = packet_radius(recs[j].w, j % 5 == 4, P.grow, pad1, pad2);
                    const bool m_cone = P.slab ? cone_pass<true>(P.k, recs[j], rm) : cone_pass<false>(P.k, recs[j], rm);
                    const bool m_ball = ball_pass(P.k.apex, ball_r, recs[j], rm);
                    member_pass += m_cone + m_ball;
                    group_bad += (m_cone && !g_cone) + (m_ball && !g_ball);
                }
            }
        }
    }
    printf("%d %ld %ld %ld %ld\n", cases, must, bad, member_pass, group_bad);
    return 0;
}
