// nrb_math.cuh — fp32 vector helpers, RNG and sRGB for the frame kernels.
//
// The frame translation units are compiled with -fmad=false: every expression below is evaluated
// as separate IEEE mul/add/sub/div/sqrt in the written order, which makes traversal, the triangle
// test, hit data and ray generation bit-identical to the reference's headers compiled for the
// host (the CPU oracle, g++ -ffp-contract=off).  Only libm calls (sinf, cosf, powf, ...) differ
// from glibc, by a few ulp.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace nrb {

struct V3 {
    float x, y, z;
};
struct V2 {
    float x, y;
};

__host__ __device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__host__ __device__ __forceinline__ V3 v3s(float s) { return V3{s, s, s}; }
__host__ __device__ __forceinline__ V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__host__ __device__ __forceinline__ V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__host__ __device__ __forceinline__ V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
__host__ __device__ __forceinline__ V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
__host__ __device__ __forceinline__ V3 operator*(float s, V3 a) { return V3{a.x * s, a.y * s, a.z * s}; }
__host__ __device__ __forceinline__ V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
__host__ __device__ __forceinline__ V3 hmul(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
// reference src/vec3.h: dot = (x*x' + y*y') + z*z', cross component = a*b - c*d
__host__ __device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__host__ __device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__host__ __device__ __forceinline__ float length(V3 a) { return sqrtf(dot(a, a)); }
__host__ __device__ __forceinline__ V3 normalize(V3 a) {
    float l = length(a);
    return l > 0.0f ? a / l : V3{0.0f, 0.0f, 0.0f};
}
__host__ __device__ __forceinline__ V3 mix3(V3 a, V3 b, float t) { return a * (1.0f - t) + b * t; }
__host__ __device__ __forceinline__ float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
__host__ __device__ __forceinline__ float clamp01(float x) { return fminf(fmaxf(x, 0.0f), 1.0f); }
__host__ __device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

__host__ __device__ __forceinline__ V3 ld3(const float* p, size_t i) { return V3{p[3 * i], p[3 * i + 1], p[3 * i + 2]}; }

// Wang hash RNG: reference src/mesh_traversal.cuh:17-34.
__host__ __device__ __forceinline__ uint32_t wang(uint32_t x) {
    x = (x ^ 61u) ^ (x >> 16);
    x *= 9u;
    x ^= x >> 4;
    x *= 0x27d4eb2du;
    x ^= x >> 15;
    return x;
}
__host__ __device__ __forceinline__ uint32_t rng_seed(int pixel, uint32_t sampleOffset, int s) {
    return wang(static_cast<uint32_t>(pixel) * 9781u + (sampleOffset + static_cast<uint32_t>(s) + 1u) * 6271u);
}
__host__ __device__ __forceinline__ float rng_next(uint32_t& st) {
    st = wang(st);
    return static_cast<float>(st & 0x00FFFFFFu) * (1.0f / 16777216.0f);
}

// sRGB OETF + clamp: reference src/cuda_renderer_neural.cu:94-113.
__device__ __forceinline__ float srgb_encode(float v) {
    v = fmaxf(0.0f, v);
    float r = v <= 0.0031308f ? 12.92f * v : 1.055f * powf(v, 1.0f / 2.4f) - 0.055f;
    return fminf(1.0f, r);
}

}  // namespace nrb
