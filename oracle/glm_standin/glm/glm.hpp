// Stand-in for the seven glm symbols the reference rasterizer uses.
//
// TEST INFRASTRUCTURE ONLY (oracle/): this header exists so that the reference's
// own CUDA sources can be compiled, where they lie under /root/reference, into
// oracle/_ref/ (the golden implementation the parity tests compare against).
// glm (https://github.com/g-truc/glm) is an un-vendored git submodule of the
// reference (submodules/diff-gaussian-rasterization/.gitmodules:1-3) and is absent
// from this image; no pinned commit is recoverable from the mount. The semantics
// restated here are glm's published ones: column-major mat3 (the 9-scalar
// constructor fills columns, m[i] is column i), mat3*mat3 evaluated per element as
// a0*b0 + a1*b1 + a2*b2 left to right, component-wise vec/scalar operators,
// dot(a,b) = a.x*b.x + a.y*b.y + a.z*b.z, length = sqrt(dot).
// Nothing in the product (3dgs-mcmc_b200/) includes this file.
#pragma once
#include <cmath>
#include <cuda_runtime.h>

#define GLM_STANDIN_FN __host__ __device__ __forceinline__

namespace glm {

struct vec3 {
    float x, y, z;
    GLM_STANDIN_FN vec3() {}
    template <typename A, typename B, typename C>
    GLM_STANDIN_FN vec3(A a, B b, C c) : x(static_cast<float>(a)), y(static_cast<float>(b)), z(static_cast<float>(c)) {}
    GLM_STANDIN_FN float& operator[](int i) { return (&x)[i]; }
    GLM_STANDIN_FN const float& operator[](int i) const { return (&x)[i]; }
    GLM_STANDIN_FN vec3& operator+=(const vec3& o) { x += o.x; y += o.y; z += o.z; return *this; }
    GLM_STANDIN_FN vec3& operator+=(float s) { x += s; y += s; z += s; return *this; }
    GLM_STANDIN_FN vec3& operator*=(float s) { x *= s; y *= s; z *= s; return *this; }
};

struct alignas(16) vec4 {
    float x, y, z, w;
    GLM_STANDIN_FN vec4() {}
    template <typename A, typename B, typename C, typename D>
    GLM_STANDIN_FN vec4(A a, B b, C c, D d) : x(static_cast<float>(a)), y(static_cast<float>(b)), z(static_cast<float>(c)), w(static_cast<float>(d)) {}
};

GLM_STANDIN_FN vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
GLM_STANDIN_FN vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
GLM_STANDIN_FN vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }
GLM_STANDIN_FN vec3 operator*(float s, const vec3& v) { return vec3(s * v.x, s * v.y, s * v.z); }
GLM_STANDIN_FN vec3 operator*(const vec3& v, float s) { return vec3(v.x * s, v.y * s, v.z * s); }
GLM_STANDIN_FN vec3 operator/(const vec3& v, float s) { return vec3(v.x / s, v.y / s, v.z / s); }

GLM_STANDIN_FN float dot(const vec3& a, const vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GLM_STANDIN_FN float length(const vec3& v) { return sqrt(dot(v, v)); }
GLM_STANDIN_FN vec3 max(const vec3& v, float s) { return vec3(fmaxf(v.x, s), fmaxf(v.y, s), fmaxf(v.z, s)); }

struct mat3 {
    vec3 c[3];  // columns
    GLM_STANDIN_FN mat3() {}
    GLM_STANDIN_FN explicit mat3(float d) { c[0] = vec3(d, 0, 0); c[1] = vec3(0, d, 0); c[2] = vec3(0, 0, d); }
    template <typename X0, typename Y0, typename Z0, typename X1, typename Y1, typename Z1, typename X2, typename Y2, typename Z2>
    GLM_STANDIN_FN mat3(X0 x0, Y0 y0, Z0 z0, X1 x1, Y1 y1, Z1 z1, X2 x2, Y2 y2, Z2 z2) {
        c[0] = vec3(x0, y0, z0); c[1] = vec3(x1, y1, z1); c[2] = vec3(x2, y2, z2);
    }
    GLM_STANDIN_FN vec3& operator[](int i) { return c[i]; }
    GLM_STANDIN_FN const vec3& operator[](int i) const { return c[i]; }
};

GLM_STANDIN_FN mat3 operator*(const mat3& a, const mat3& b) {
    mat3 r;
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 3; ++i)
            r[j][i] = a[0][i] * b[j][0] + a[1][i] * b[j][1] + a[2][i] * b[j][2];
    return r;
}
GLM_STANDIN_FN vec3 operator*(const mat3& m, const vec3& v) {
    return vec3(m[0][0] * v.x + m[1][0] * v.y + m[2][0] * v.z,
                m[0][1] * v.x + m[1][1] * v.y + m[2][1] * v.z,
                m[0][2] * v.x + m[1][2] * v.y + m[2][2] * v.z);
}
GLM_STANDIN_FN mat3 operator*(float s, const mat3& m) {
    mat3 r; r[0] = m[0] * s; r[1] = m[1] * s; r[2] = m[2] * s; return r;
}
GLM_STANDIN_FN mat3 transpose(const mat3& m) {
    return mat3(m[0][0], m[1][0], m[2][0], m[0][1], m[1][1], m[2][1], m[0][2], m[1][2], m[2][2]);
}

}  // namespace glm
