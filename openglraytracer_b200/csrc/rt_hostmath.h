/*
 * rt_hostmath.h — FP32 host prologue math: the pixel-invariant and
 * ray-invariant matrices the shader rebuilds per pixel / per intersection test
 * (glsl:366-383 camera, glsl:650-652 + 718 per box), computed ONCE per frame /
 * per object here.
 *
 * Matrices are flat float[16], column-major like GLSL: element (row r, col c)
 * is m[4*c + r].  Operation order is fixed (sums left to right over k; adjugate
 * inverse from 2x2 sub-determinants) and the file is compiled with
 * -ffp-contract=off, so results do not depend on the compiler's FMA choices.
 */
#ifndef RT_HOSTMATH_H
#define RT_HOSTMATH_H

#include <math.h>
#include <string.h>

/* The device prologue (rt_prologue.cu, compiled -fmad=false) runs the same functions on the GPU. */
#ifdef __CUDACC__
#define RTHM_FN __host__ __device__ inline
#else
#define RTHM_FN inline
#endif

namespace rthm {

constexpr float kPi = 3.14159265358f;          /* glsl:13 */
constexpr float kDegToRad = kPi / 180.0f;      /* glsl:17 */

RTHM_FN float& at(float* m, int r, int c) { return m[4 * c + r]; }
RTHM_FN float at(const float* m, int r, int c) { return m[4 * c + r]; }

RTHM_FN void identity(float* m) {
    for (int i = 0; i < 16; i++) m[i] = 0.0f;
    m[0] = m[5] = m[10] = m[15] = 1.0f;
}

/* out = a * b; out may not alias a or b */
RTHM_FN void mul(const float* a, const float* b, float* out) {
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++) {
            float acc = at(a, r, 0) * at(b, 0, c);
            acc = acc + at(a, r, 1) * at(b, 1, c);
            acc = acc + at(a, r, 2) * at(b, 2, c);
            acc = acc + at(a, r, 3) * at(b, 3, c);
            at(out, r, c) = acc;
        }
}

/* glsl:444-486 */
RTHM_FN void rot_x(float deg, float* m) {
    float c = cosf(kDegToRad * deg), s = sinf(kDegToRad * deg);
    identity(m);
    at(m, 1, 1) = c; at(m, 2, 1) = s; at(m, 1, 2) = -s; at(m, 2, 2) = c;
}
RTHM_FN void rot_y(float deg, float* m) {
    float c = cosf(kDegToRad * deg), s = sinf(kDegToRad * deg);
    identity(m);
    at(m, 0, 0) = c; at(m, 2, 0) = -s; at(m, 0, 2) = s; at(m, 2, 2) = c;
}
RTHM_FN void rot_z(float deg, float* m) {
    float c = cosf(kDegToRad * deg), s = sinf(kDegToRad * deg);
    identity(m);
    at(m, 0, 0) = c; at(m, 1, 0) = s; at(m, 0, 1) = -s; at(m, 1, 1) = c;
}

/* glsl:492-503: I * Rz(yaw) * Rx(pitch) * Ry(roll), angles = (pitch, yaw, roll) */
RTHM_FN void rotation(const float* angles, float* out) {
    float i4[16], rz[16], rx[16], ry[16], t0[16], t1[16];
    identity(i4);
    rot_z(angles[1], rz); rot_x(angles[0], rx); rot_y(angles[2], ry);
    mul(i4, rz, t0);
    mul(t0, rx, t1);
    mul(t1, ry, out);
}

/* glsl:529-532: T(position) * R(angles) */
RTHM_FN void transform(const float* position, const float* angles, float* out) {
    float t[16], r[16];
    identity(t);
    at(t, 0, 3) = position[0]; at(t, 1, 3) = position[1]; at(t, 2, 3) = position[2]; at(t, 3, 3) = 1.0f;
    rotation(angles, r);
    mul(t, r, out);
}

/* inverse(mat4): adjugate from 2x2 sub-determinants, times 1/det */
RTHM_FN void inverse4(const float* m, float* out) {
#define A(i, j) at(m, i, j)
    float s0 = A(0,0) * A(1,1) - A(1,0) * A(0,1);
    float s1 = A(0,0) * A(1,2) - A(1,0) * A(0,2);
    float s2 = A(0,0) * A(1,3) - A(1,0) * A(0,3);
    float s3 = A(0,1) * A(1,2) - A(1,1) * A(0,2);
    float s4 = A(0,1) * A(1,3) - A(1,1) * A(0,3);
    float s5 = A(0,2) * A(1,3) - A(1,2) * A(0,3);
    float c5 = A(2,2) * A(3,3) - A(3,2) * A(2,3);
    float c4 = A(2,1) * A(3,3) - A(3,1) * A(2,3);
    float c3 = A(2,1) * A(3,2) - A(3,1) * A(2,2);
    float c2 = A(2,0) * A(3,3) - A(3,0) * A(2,3);
    float c1 = A(2,0) * A(3,2) - A(3,0) * A(2,2);
    float c0 = A(2,0) * A(3,1) - A(3,0) * A(2,1);
    float det = ((((s0 * c5 - s1 * c4) + s2 * c3) + s3 * c2) - s4 * c1) + s5 * c0;
    float id = 1.0f / det;
    at(out, 0, 0) = ((A(1,1) * c5 - A(1,2) * c4) + A(1,3) * c3) * id;
    at(out, 0, 1) = ((-A(0,1) * c5 + A(0,2) * c4) - A(0,3) * c3) * id;
    at(out, 0, 2) = ((A(3,1) * s5 - A(3,2) * s4) + A(3,3) * s3) * id;
    at(out, 0, 3) = ((-A(2,1) * s5 + A(2,2) * s4) - A(2,3) * s3) * id;
    at(out, 1, 0) = ((-A(1,0) * c5 + A(1,2) * c2) - A(1,3) * c1) * id;
    at(out, 1, 1) = ((A(0,0) * c5 - A(0,2) * c2) + A(0,3) * c1) * id;
    at(out, 1, 2) = ((-A(3,0) * s5 + A(3,2) * s2) - A(3,3) * s1) * id;
    at(out, 1, 3) = ((A(2,0) * s5 - A(2,2) * s2) + A(2,3) * s1) * id;
    at(out, 2, 0) = ((A(1,0) * c4 - A(1,1) * c2) + A(1,3) * c0) * id;
    at(out, 2, 1) = ((-A(0,0) * c4 + A(0,1) * c2) - A(0,3) * c0) * id;
    at(out, 2, 2) = ((A(3,0) * s4 - A(3,1) * s2) + A(3,3) * s0) * id;
    at(out, 2, 3) = ((-A(2,0) * s4 + A(2,1) * s2) - A(2,3) * s0) * id;
    at(out, 3, 0) = ((-A(1,0) * c3 + A(1,1) * c1) - A(1,2) * c0) * id;
    at(out, 3, 1) = ((A(0,0) * c3 - A(0,1) * c1) + A(0,2) * c0) * id;
    at(out, 3, 2) = ((-A(3,0) * s3 + A(3,1) * s1) - A(3,2) * s0) * id;
    at(out, 3, 3) = ((A(2,0) * s3 - A(2,1) * s1) + A(2,2) * s0) * id;
#undef A
}

/* transpose(inverse(mat3(m))) as 9 floats, column-major n[3*c + r] (glsl:718) */
RTHM_FN void normal_matrix(const float* m, float* n) {
#define A(i, j) at(m, i, j)
    float k00 = A(1,1) * A(2,2) - A(1,2) * A(2,1);
    float k01 = A(1,2) * A(2,0) - A(1,0) * A(2,2);
    float k02 = A(1,0) * A(2,1) - A(1,1) * A(2,0);
    float det = (A(0,0) * k00 + A(0,1) * k01) + A(0,2) * k02;
    float id = 1.0f / det;
    float inv[3][3];   /* inv[r][c] */
    inv[0][0] = k00 * id;
    inv[0][1] = (A(0,2) * A(2,1) - A(0,1) * A(2,2)) * id;
    inv[0][2] = (A(0,1) * A(1,2) - A(0,2) * A(1,1)) * id;
    inv[1][0] = k01 * id;
    inv[1][1] = (A(0,0) * A(2,2) - A(0,2) * A(2,0)) * id;
    inv[1][2] = (A(0,2) * A(1,0) - A(0,0) * A(1,2)) * id;
    inv[2][0] = k02 * id;
    inv[2][1] = (A(0,1) * A(2,0) - A(0,0) * A(2,1)) * id;
    inv[2][2] = (A(0,0) * A(1,1) - A(0,1) * A(1,0)) * id;
#undef A
    /* transpose: n(r,c) = inv(c,r) */
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++) n[3 * c + r] = inv[c][r];
}

/* glsl:411-426 */
RTHM_FN void projection(float v_fov, float aspect, float near_plane, float far_plane, float* m) {
    float q = 1.0f / tanf(kDegToRad * 0.5f * v_fov);
    float A = q / aspect;
    float B = (near_plane + far_plane) / (near_plane - far_plane);
    float C = (2.0f * near_plane * far_plane) / (near_plane - far_plane);
    for (int i = 0; i < 16; i++) m[i] = 0.0f;
    at(m, 0, 0) = A;
    at(m, 1, 1) = q;
    at(m, 2, 2) = B;
    at(m, 3, 2) = -1.0f;   /* result[2][3] in GLSL column-major indexing */
    at(m, 2, 3) = C;       /* result[3][2] */
}

/* glsl:538-545: inverse(T*R * Rx(90)) */
RTHM_FN void view(const float* position, const float* angles, float* out) {
    float tr[16], flip[16], prod[16];
    transform(position, angles, tr);
    rot_x(90.0f, flip);
    mul(tr, flip, prod);
    inverse4(prod, out);
}

/* glsl:383: inverse(proj * view) */
RTHM_FN void inverse_proj_view(const float* position, const float* angles, float v_fov, float aspect,
                              float near_plane, float far_plane, float* out) {
    float p[16], v[16], pv[16];
    projection(v_fov, aspect, near_plane, far_plane, p);
    view(position, angles, v);
    mul(p, v, pv);
    inverse4(pv, out);
}

}  // namespace rthm
#endif
