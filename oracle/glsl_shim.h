/*
 * glsl_shim.h — the GLSL 4.30 value types and built-ins the shader uses, in FP32,
 * with every implementation-defined behaviour PINNED in writing:
 *
 *   dot(a,b)        = ((a.x*b.x + a.y*b.y) + a.z*b.z) [+ a.w*b.w], left to right
 *   mat*vec, mat*mat: per output element, sum over k = 0..3 left to right
 *   normalize(v)    = v / sqrt(dot(v,v))            (IEEE sqrt, IEEE divide per component)
 *   reflect(I,N)    = I - 2*dot(N,I)*N
 *   refract(I,N,e)  = k<0 ? 0 : e*I - (e*dot(N,I) + sqrt(k))*N,  k = 1 - e*e*(1 - dot(N,I)^2)
 *   mix(x,y,a)      = x*(1-a) + y*a
 *   mod(x,y)        = x - y*floor(x/y)
 *   min/max         = fminf/fmaxf (return the non-NaN operand)
 *   inverse(mat4)   = adjugate from 2x2 sub-determinants s0..s5,c0..c5, times 1/det
 *   inverse(mat3)   = cofactors times 1/det
 *   float -> UNORM8 = (uint8) rintf(fminf(fmaxf(f,0),1) * 255.0f), NaN -> 0
 *   sin/cos/tan/pow = glibc sinf/cosf/tanf/powf
 *
 * Compiled with -ffp-contract=off so no FMA is formed.  TEST INFRASTRUCTURE
 * (see rt_oracle.h); not part of the product.
 */
#ifndef GLSL_SHIM_H
#define GLSL_SHIM_H

#include <cmath>
#include <cstdint>

namespace glsl {

struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};
struct vec4 {
    float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec4(const vec3& v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
    vec3 xyz() const { return vec3(x, y, z); }
    vec3 rgb() const { return vec3(x, y, z); }
};

inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline bool operator==(vec3 a, vec3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
inline bool operator!=(vec3 a, vec3 b) { return !(a == b); }

inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(vec4 a, vec4 b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline vec4 operator*(vec4 a, vec4 b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator/(vec4 a, float s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }

inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float gmin(float a, float b) { return fminf(a, b); }
inline float gmax(float a, float b) { return fmaxf(a, b); }
inline vec3 gmin(vec3 a, vec3 b) { return vec3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
inline vec3 gmax(vec3 a, vec3 b) { return vec3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
inline vec3 normalize(vec3 v) { float l = sqrtf(dot(v, v)); return vec3(v.x / l, v.y / l, v.z / l); }
inline vec3 reflect(vec3 I, vec3 N) { float d2 = 2.0f * dot(N, I); return I - d2 * N; }
inline vec3 refract(vec3 I, vec3 N, float eta) {
    float ni = dot(N, I);
    float k = 1.0f - eta * eta * (1.0f - ni * ni);
    if (k < 0.0f) return vec3(0.0f);
    return eta * I - (eta * ni + sqrtf(k)) * N;
}
inline vec3 mix(vec3 x, vec3 y, float a) { float ia = 1.0f - a; return x * ia + y * a; }
inline float mod(float x, float y) { return x - y * floorf(x / y); }

/* Column-major like GLSL: m.c[col][row]. */
struct mat4 {
    float c[4][4];
    mat4() { for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) c[i][j] = 0.0f; }
    explicit mat4(float d) { for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) c[i][j] = (i == j) ? d : 0.0f; }
};
struct mat3 {
    float c[3][3];
    mat3() { for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) c[i][j] = 0.0f; }
    explicit mat3(const mat4& m) { for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) c[i][j] = m.c[i][j]; }
};

inline mat4 operator*(const mat4& a, const mat4& b) {
    mat4 r;
    for (int col = 0; col < 4; col++)
        for (int row = 0; row < 4; row++)
            r.c[col][row] = ((a.c[0][row] * b.c[col][0] + a.c[1][row] * b.c[col][1]) + a.c[2][row] * b.c[col][2]) + a.c[3][row] * b.c[col][3];
    return r;
}
inline vec4 operator*(const mat4& m, vec4 v) {
    vec4 r;
    for (int row = 0; row < 4; row++)
        r[row] = ((m.c[0][row] * v.x + m.c[1][row] * v.y) + m.c[2][row] * v.z) + m.c[3][row] * v.w;
    return r;
}
inline vec3 operator*(const mat3& m, vec3 v) {
    vec3 r;
    for (int row = 0; row < 3; row++)
        r[row] = (m.c[0][row] * v.x + m.c[1][row] * v.y) + m.c[2][row] * v.z;
    return r;
}
inline mat3 transpose(const mat3& m) {
    mat3 r;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) r.c[i][j] = m.c[j][i];
    return r;
}

/* inverse(mat4): a(i,j) = element at row i, column j. */
inline mat4 inverse(const mat4& m) {
#define A(i, j) m.c[j][i]
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
    mat4 r;
#define B(i, j) r.c[j][i]
    B(0,0) = ((A(1,1) * c5 - A(1,2) * c4) + A(1,3) * c3) * id;
    B(0,1) = ((-A(0,1) * c5 + A(0,2) * c4) - A(0,3) * c3) * id;
    B(0,2) = ((A(3,1) * s5 - A(3,2) * s4) + A(3,3) * s3) * id;
    B(0,3) = ((-A(2,1) * s5 + A(2,2) * s4) - A(2,3) * s3) * id;
    B(1,0) = ((-A(1,0) * c5 + A(1,2) * c2) - A(1,3) * c1) * id;
    B(1,1) = ((A(0,0) * c5 - A(0,2) * c2) + A(0,3) * c1) * id;
    B(1,2) = ((-A(3,0) * s5 + A(3,2) * s2) - A(3,3) * s1) * id;
    B(1,3) = ((A(2,0) * s5 - A(2,2) * s2) + A(2,3) * s1) * id;
    B(2,0) = ((A(1,0) * c4 - A(1,1) * c2) + A(1,3) * c0) * id;
    B(2,1) = ((-A(0,0) * c4 + A(0,1) * c2) - A(0,3) * c0) * id;
    B(2,2) = ((A(3,0) * s4 - A(3,1) * s2) + A(3,3) * s0) * id;
    B(2,3) = ((-A(2,0) * s4 + A(2,1) * s2) - A(2,3) * s0) * id;
    B(3,0) = ((-A(1,0) * c3 + A(1,1) * c1) - A(1,2) * c0) * id;
    B(3,1) = ((A(0,0) * c3 - A(0,1) * c1) + A(0,2) * c0) * id;
    B(3,2) = ((-A(3,0) * s3 + A(3,1) * s1) - A(3,2) * s0) * id;
    B(3,3) = ((A(2,0) * s3 - A(2,1) * s1) + A(2,2) * s0) * id;
#undef A
#undef B
    return r;
}

inline mat3 inverse(const mat3& m) {
#define A(i, j) m.c[j][i]
    float k00 = A(1,1) * A(2,2) - A(1,2) * A(2,1);
    float k01 = A(1,2) * A(2,0) - A(1,0) * A(2,2);
    float k02 = A(1,0) * A(2,1) - A(1,1) * A(2,0);
    float det = (A(0,0) * k00 + A(0,1) * k01) + A(0,2) * k02;
    float id = 1.0f / det;
    mat3 r;
#define B(i, j) r.c[j][i]
    B(0,0) = k00 * id;
    B(0,1) = (A(0,2) * A(2,1) - A(0,1) * A(2,2)) * id;
    B(0,2) = (A(0,1) * A(1,2) - A(0,2) * A(1,1)) * id;
    B(1,0) = k01 * id;
    B(1,1) = (A(0,0) * A(2,2) - A(0,2) * A(2,0)) * id;
    B(1,2) = (A(0,2) * A(1,0) - A(0,0) * A(1,2)) * id;
    B(2,0) = k02 * id;
    B(2,1) = (A(0,1) * A(2,0) - A(0,0) * A(2,1)) * id;
    B(2,2) = (A(0,0) * A(1,1) - A(0,1) * A(1,0)) * id;
#undef A
#undef B
    return r;
}

inline uint8_t unorm8(float f) {
    if (f != f) return 0;
    float c = fminf(fmaxf(f, 0.0f), 1.0f);
    return (uint8_t)rintf(c * 255.0f);
}

}  // namespace glsl
#endif
