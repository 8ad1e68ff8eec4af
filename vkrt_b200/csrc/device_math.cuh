// device_math.cuh — FP32 vector helpers with a fixed evaluation order.
//
// The functions are `__host__ __device__` so that the kernel LOGIC can also be compiled by g++ into
// the host emulation harness under tests/emu/ (a no-GPU unit-test aid; it is not part of the
// product library and nothing in libvkrt_b200.so calls it).
#pragma once

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define VKRT_HD __host__ __device__ __forceinline__
#define VKRT_D __device__ __forceinline__
#else
#define VKRT_HD inline
#define VKRT_D inline
#endif

namespace vkrt
{

struct F3
{
    float x, y, z;
};

VKRT_HD F3 f3(float x, float y, float z)
{
    F3 r;
    r.x = x, r.y = y, r.z = z;
    return r;
}
VKRT_HD F3 operator+(F3 a, F3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
VKRT_HD F3 operator-(F3 a, F3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
VKRT_HD F3 operator*(F3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
VKRT_HD F3 operator*(F3 a, F3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
VKRT_HD F3 operator/(F3 a, float s) { return f3(a.x / s, a.y / s, a.z / s); }
// dot = (x*x' + y*y') + z*z'
VKRT_HD float dot3(F3 a, F3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
VKRT_HD float length3(F3 a) { return sqrtf(dot3(a, a)); }
// GLSL normalize: v / length(v)
VKRT_HD F3 normalize3(F3 a) { return a / length3(a); }
// GLSL cross
VKRT_HD F3 cross3(F3 a, F3 b) { return f3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
// GLSL clamp(x,0,1) as min(max(x,0),1); NaN -> 0 (fmaxf drops the NaN operand on both CPU and GPU)
VKRT_HD float clamp01(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }
VKRT_HD float comp3(F3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
VKRT_HD bool finite1(float x)
{
    // true for anything but inf/NaN; written with integer ops so -ffast-math style folding cannot break it
    uint32_t u;
#if defined(__CUDA_ARCH__)
    u = __float_as_uint(x);
#else
    memcpy(&u, &x, 4);
#endif
    return (u & 0x7F800000u) != 0x7F800000u;
}
VKRT_HD bool finite3(F3 v) { return finite1(v.x) && finite1(v.y) && finite1(v.z); }

VKRT_HD float uintAsFloat(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}
VKRT_HD uint32_t floatAsUint(float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
#endif
}

// Explicit fused multiply-add, used only where the result must be conservative, never on the parity path.
VKRT_HD float fmaExplicit(float a, float b, float c)
{
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}

// column-major mat4 * (x,y,z,w), rows 0..2:  ((c0*x + c1*y) + c2*z) + c3*w
VKRT_HD F3 mat4MulVec(const float* m, float x, float y, float z, float w)
{
    F3 r;
    r.x = ((m[0] * x + m[4] * y) + m[8] * z) + m[12] * w;
    r.y = ((m[1] * x + m[5] * y) + m[9] * z) + m[13] * w;
    r.z = ((m[2] * x + m[6] * y) + m[10] * z) + m[14] * w;
    return r;
}

// row-major 3x4 affine applied to a point / a direction
VKRT_HD F3 affinePoint(const float* m, F3 p)
{
    F3 r;
    r.x = ((m[0] * p.x + m[1] * p.y) + m[2] * p.z) + m[3];
    r.y = ((m[4] * p.x + m[5] * p.y) + m[6] * p.z) + m[7];
    r.z = ((m[8] * p.x + m[9] * p.y) + m[10] * p.z) + m[11];
    return r;
}
VKRT_HD F3 affineDir(const float* m, F3 d)
{
    F3 r;
    r.x = (m[0] * d.x + m[1] * d.y) + m[2] * d.z;
    r.y = (m[4] * d.x + m[5] * d.y) + m[6] * d.z;
    r.z = (m[8] * d.x + m[9] * d.y) + m[10] * d.z;
    return r;
}

// Inverse of a row-major 3x4 affine transform: double cofactors, rounded to float once.
inline void affineInverseHost(const float m[12], float inv[12])
{
    const double a = m[0], b = m[1], c = m[2], tx = m[3];
    const double d = m[4], e = m[5], f = m[6], ty = m[7];
    const double g = m[8], h = m[9], i = m[10], tz = m[11];
    const double c00 = e * i - f * h;
    const double c01 = f * g - d * i;
    const double c02 = d * h - e * g;
    const double det = (a * c00 + b * c01) + c * c02;
    const double r00 = c00 / det, r01 = (c * h - b * i) / det, r02 = (b * f - c * e) / det;
    const double r10 = c01 / det, r11 = (a * i - c * g) / det, r12 = (c * d - a * f) / det;
    const double r20 = c02 / det, r21 = (b * g - a * h) / det, r22 = (a * e - b * d) / det;
    inv[0] = (float)r00, inv[1] = (float)r01, inv[2] = (float)r02;
    inv[4] = (float)r10, inv[5] = (float)r11, inv[6] = (float)r12;
    inv[8] = (float)r20, inv[9] = (float)r21, inv[10] = (float)r22;
    inv[3] = (float)(-((r00 * tx + r01 * ty) + r02 * tz));
    inv[7] = (float)(-((r10 * tx + r11 * ty) + r12 * tz));
    inv[11] = (float)(-((r20 * tx + r21 * ty) + r22 * tz));
}

} // namespace vkrt
