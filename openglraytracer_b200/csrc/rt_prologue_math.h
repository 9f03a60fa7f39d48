/*
 * rt_prologue_math.h — the per-object and per-(object, light) prologue, written once and run either on
 * the host (rt_host.cpp) or on the GPU (rt_prologue.cu): what the shader rebuilds per intersection
 * test (glsl:650-652 local_to_world + inverse, glsl:718 normal matrix) hoisted to once per object,
 * plus the records of the large-scene loops.
 *
 * The only difference between the two runs is the sinf / cosf implementation (glibc vs CUDA, both
 * within 1-2 ulp): the device prologue is therefore used by the fast kernel only; the strict kernel
 * (op-for-op parity with the oracle) always takes its records from the host run.
 */
#ifndef RT_PROLOGUE_MATH_H
#define RT_PROLOGUE_MATH_H

#include "../../include/rt_b200.h"
#include "rt_device.h"
#include "rt_hostmath.h"

namespace rtpm {

/* glsl:749 / 759: box iff box != null_box; else sphere iff radius != -1; else skipped */
RTHM_FN int object_kind(const rt_object& o) {
    const bool null_box = o.box_mins[0] == 0.0f && o.box_mins[1] == 0.0f && o.box_mins[2] == 0.0f &&
                          o.box_maxs[0] == 0.0f && o.box_maxs[1] == 0.0f && o.box_maxs[2] == 0.0f;
    if (!null_box) return RT_KIND_BOX;
    if (o.sphere_radius != -1.0f) return RT_KIND_SPHERE;
    return RT_KIND_SKIP;
}

/*
 * Records of one object.  cold: always.  Boxes: fast12 (rows of diag(1/h) * (w2l - centre)), strict20
 * (w2l rows + mins + maxs, may be NULL), cull4 = (world centre, UN-padded circumradius, < 0 for a
 * zero-thickness box that can never satisfy t_near < t_far, glsl:680).  Spheres: sph4 = (centre, r*r).
 */
RTHM_FN void object_records(const rt_object& o, int kind, int hot_index, RtCold* cp,
                            float* fast12, float* strict20, float* cull4, float* sph4) {
    RtCold& c = *cp;
    for (int k = 0; k < 4; k++) {
        c.ambient[k] = o.material.ambient[k]; c.diffuse[k] = o.material.diffuse[k];
        c.specular[k] = o.material.specular[k]; c.emissive[k] = o.material.emissive[k];
    }
    c.shininess = o.material.shininess;
    c.reflectivity = o.material.reflectivity;
    c.transparency = o.material.transparency;
    c.refraction_index = o.material.refraction_index;
    for (int k = 0; k < 3; k++) c.position[k] = o.position[k];
    c.inv_ior = 1.0f / o.material.refraction_index;
    c.ior_back = 1.0f / c.inv_ior;
    for (int k = 0; k < 4; k++) c.unused_[k] = 0.0f;
    c.kind = kind;
    c.hot_index = hot_index;
    for (int k = 0; k < 12; k++) { c.w2l[k] = 0.0f; c.l2w[k] = 0.0f; }
    for (int k = 0; k < 9; k++) c.nmat[k] = 0.0f;
    if (kind == RT_KIND_BOX) {
        float l2w[16], w2l[16];
        rthm::transform(o.position, o.angles, l2w);     // glsl:650
        rthm::inverse4(l2w, w2l);                       // glsl:652
        rthm::normal_matrix(l2w, c.nmat);               // glsl:718
        for (int r = 0; r < 3; r++)
            for (int k = 0; k < 4; k++) {
                c.w2l[4 * r + k] = rthm::at(w2l, r, k);
                c.l2w[4 * r + k] = rthm::at(l2w, r, k);
            }
        const bool degenerate = o.box_mins[0] == o.box_maxs[0] || o.box_mins[1] == o.box_maxs[1] ||
                                o.box_mins[2] == o.box_maxs[2];
        if (strict20) {
            for (int k = 0; k < 12; k++) strict20[k] = c.w2l[k];
            for (int k = 0; k < 3; k++) { strict20[12 + k] = o.box_mins[k]; strict20[15 + k] = o.box_maxs[k]; }
            strict20[18] = 0.0f; strict20[19] = 0.0f;
        }
        double ctr[3], r2 = 0.0;
        for (int r = 0; r < 3; r++) {
            const double h = 0.5 * ((double)o.box_maxs[r] - (double)o.box_mins[r]);
            ctr[r] = 0.5 * ((double)o.box_maxs[r] + (double)o.box_mins[r]);
            r2 += h * h;
            if (degenerate) {
                // unreachable slab: origin far outside, direction row zero -> always a miss
                fast12[4 * r + 0] = 0.0f; fast12[4 * r + 1] = 0.0f; fast12[4 * r + 2] = 0.0f; fast12[4 * r + 3] = 4.0f;
            } else {
                const double ih = 1.0 / h;
                for (int k = 0; k < 3; k++) fast12[4 * r + k] = (float)((double)c.w2l[4 * r + k] * ih);
                fast12[4 * r + 3] = (float)(((double)c.w2l[4 * r + 3] - ctr[r]) * ih);
            }
        }
        // bounding sphere in world space (T*R is rigid: the circumradius is |half extent|)
        for (int r = 0; r < 3; r++)
            cull4[r] = (float)((double)c.l2w[4 * r + 0] * ctr[0] + (double)c.l2w[4 * r + 1] * ctr[1] +
                               (double)c.l2w[4 * r + 2] * ctr[2] + (double)c.l2w[4 * r + 3]);
        cull4[3] = degenerate ? -1.0f : (float)sqrt(r2);
    } else if (kind == RT_KIND_SPHERE) {
        sph4[0] = o.position[0]; sph4[1] = o.position[1]; sph4[2] = o.position[2];
        sph4[3] = o.sphere_radius * o.sphere_radius;    // glsl:588 radius * radius
    }
}

/* |centre| + radius of a hot object: its contribution to the scene extent E */
RTHM_FN double reach(const float* c3, double radius) {
    return sqrt((double)c3[0] * c3[0] + (double)c3[1] * c3[1] + (double)c3[2] * c3[2]) + radius;
}

/*
 * Padding.  Ray origins, lights and centres have coordinates up to ~E, so
 * a difference of two of them carries an absolute FP32 rounding error up to ~E * 2^-23; bounding radii
 * are padded by 5e-4 relative + 2e-6 * E absolute.  Light records grow the radius by a further 0.0101
 * (the shadow ray's line is the line through the light shifted by 0.01 n, glsl:808-809), and sphere
 * records by the cancellation noise of the shader's own quadratic (glsl:586-591), 4e-6 * E^2 in R^2.
 */
RTHM_FN float padded_radius2(double radius, double extent) {
    if (radius < 0.0) return -1e30f;
    const double rp = radius * 1.0005 + 2e-6 * extent;
    return (float)(rp * rp * (1.0 + 1e-6));
}

RTHM_FN void light_record(const float* centre, double radius, bool is_box, const float* light_pos, double extent,
                          float* rec4) {
    if (radius < 0.0) { rec4[0] = 0.0f; rec4[1] = 0.0f; rec4[2] = 0.0f; rec4[3] = 1e30f; return; }   // never passes
    const double rp = radius * 1.0005 + 0.0101 + 2e-6 * extent;
    const double rp2 = rp * rp + (is_box ? 0.0 : 4e-6 * extent * extent);
    double oc2 = 0.0;
    for (int r = 0; r < 3; r++) {
        const double oc = (double)light_pos[r] - (double)centre[r];
        rec4[r] = (float)oc;
        oc2 += oc * oc;
    }
    const double qc = oc2 - rp2;
    rec4[3] = (float)(qc - fabs(qc) * 1e-6 - 1e-12);   // rounded toward "pass"
}

}  // namespace rtpm
#endif
