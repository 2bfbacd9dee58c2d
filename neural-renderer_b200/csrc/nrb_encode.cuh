// nrb_encode.cuh — device arithmetic of the input encodings, shared by both query kernels.
// Bit-faithful to tiny-cuda-nn (julcst fork @2a4340e): kernel_grid (encodings/grid.h:48-169;
// pos_fract / grid_index / coherent_prime_hash: common_device.h:768-791, 848-895, 1017-1029),
// sh_enc (common_device.h:476-502), LeakyReLU in fp16 (common_device.h:123-129).
#pragma once

#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "nrb_neural.cuh"

namespace nrb {

// fp32 accumulator -> fp16, LeakyReLU evaluated in fp16 as tiny-cuda-nn does.
__device__ __forceinline__ __half leaky_h(float v) {
    const __half h = __float2half_rn(v);
    return __hgt(h, __float2half_rn(0.0f)) ? h : __hmul(h, __float2half_rn(0.01f));
}

// Two accumulators -> packed fp16 pair with the fp16 LeakyReLU: max(h, 0.01 * h) in fp16 equals
// tiny-cuda-nn's h > 0 ? h : h * 0.01 (the product is rounded to fp16 in both) for every non-NaN h.
__device__ __forceinline__ uint32_t leaky_pack2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    const __half2 r = __hmax2(h, __hmul2(h, __float2half2_rn(0.01f)));
    return *reinterpret_cast<const uint32_t*>(&r);
}

// grid_index (common_device.h:848-884) with the coherent prime hash.
__device__ __forceinline__ uint32_t level_index(bool dense, bool pow2, uint32_t size, uint32_t res, uint32_t x,
                                                uint32_t y, uint32_t z) {
    uint32_t idx = dense ? x + y * res + z * res * res : (x ^ (y * 2654435761u) ^ (z * 805459861u));
    return pow2 ? (idx & (size - 1u)) : (idx % size);
}

// One (row, point, level) item: 8 corner gathers, fp16 FMA chain in tiny-cuda-nn's corner order.
template <int F>
__device__ __forceinline__ void encode_item(const NetDev& net, int level, const float* x, __half* dst) {
    const float scale = net.levelScale[level];
    const uint32_t res = net.levelRes[level], size = net.levelSize[level];
    const bool dense = net.levelDense[level], pow2 = net.levelPow2[level];
    const __half* table = net.grid + static_cast<size_t>(net.levelOffset[level]) * F;
    float w[3];
    uint32_t g[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float p = fmaf(scale, x[d], 0.5f);
        const float fl = floorf(p);
        g[d] = static_cast<uint32_t>(static_cast<int>(fl));
        w[d] = p - fl;
    }
    // issue all 8 gathers first (independent loads), then run the dependent fp16 chain
    uint32_t raw[8][F / 2];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const uint32_t cx = g[0] + (c & 1), cy = g[1] + ((c >> 1) & 1), cz = g[2] + ((c >> 2) & 1);
        const uint32_t idx = level_index(dense, pow2, size, res, cx, cy, cz);
        if (F == 4) {
            const uint2 v = __ldg(reinterpret_cast<const uint2*>(table + static_cast<size_t>(idx) * 4));
            raw[c][0] = v.x;
            raw[c][F / 2 - 1] = v.y;
        } else {
            raw[c][0] = __ldg(reinterpret_cast<const uint32_t*>(table + static_cast<size_t>(idx) * 2));
        }
    }
    __half2 acc[F / 2];
#pragma unroll
    for (int k = 0; k < F / 2; ++k) acc[k] = __float2half2_rn(0.0f);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        float wt = 1.0f;
#pragma unroll
        for (int d = 0; d < 3; ++d) wt *= (c & (1 << d)) ? w[d] : 1.0f - w[d];
        const __half2 wh = __half2half2(__float2half_rn(wt));
#pragma unroll
        for (int k = 0; k < F / 2; ++k) acc[k] = __hfma2(wh, *reinterpret_cast<const __half2*>(&raw[c][k]), acc[k]);
    }
#pragma unroll
    for (int k = 0; k < F / 2; ++k) *reinterpret_cast<__half2*>(dst + 2 * k) = acc[k];
}

template <int F>
__device__ __forceinline__ void encode_item(const NetDev& net, int level, float x0, float x1, float x2, __half* dst) {
    const float x[3] = {x0, x1, x2};
    encode_item<F>(net, level, x, dst);
}

// Lane-pair variant of encode_item.  The L1/TEX unit pays ~2 cycles per distinct 128-byte line a
// warp-wide load touches; with one item per lane a gather instruction touches 32 lines.  Here the
// two lanes of a pair (2i, 2i+1) split the x-corners of the pair's two items: in every load
// instruction lane 2i fetches corner (x, y', z') and lane 2i+1 the adjacent (x+1, y', z') of the SAME
// item - neighbouring table entries, one line - and the halves are exchanged with shuffles so
// that each lane ends up with all 8 corners of its own item.  The fp16 accumulation chain is the
// same corner order as tiny-cuda-nn's, so results are bit-identical to encode_item.
// Both lanes of a pair must call this with items of the same level.
template <int F>
__device__ __forceinline__ void encode_item_paired(const NetDev& net, int level, const float* x, __half* dst) {
    const float scale = net.levelScale[level];
    const uint32_t res = net.levelRes[level], size = net.levelSize[level];
    const bool dense = net.levelDense[level], pow2 = net.levelPow2[level];
    const __half* table = net.grid + static_cast<size_t>(net.levelOffset[level]) * F;
    const unsigned full = 0xffffffffu;
    const int odd = threadIdx.x & 1;
    float w[3];
    uint32_t g[3], gp[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float p = fmaf(scale, x[d], 0.5f);
        const float fl = floorf(p);
        g[d] = static_cast<uint32_t>(static_cast<int>(fl));
        w[d] = p - fl;
        gp[d] = __shfl_xor_sync(full, g[d], 1);  // the partner lane's cell
    }
    // item A = the even lane's, item B = the odd lane's; this lane loads the dx = odd corners of both
    uint32_t rawA[4][F / 2], rawB[4][F / 2];
#pragma unroll
    for (int it = 0; it < 2; ++it) {
        const bool mine = (it == odd);  // it == 0 -> item A (even lane's own)
        const uint32_t bx = (mine ? g[0] : gp[0]) + odd, by = mine ? g[1] : gp[1], bz = mine ? g[2] : gp[2];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const uint32_t idx = level_index(dense, pow2, size, res, bx, by + (c & 1), bz + (c >> 1));
            if (F == 4) {
                const uint2 v = (net.debug & 1) ? make_uint2(idx, idx) : __ldg(reinterpret_cast<const uint2*>(table + static_cast<size_t>(idx) * 4));
                if (it == 0) { rawA[c][0] = v.x; rawA[c][F / 2 - 1] = v.y; }
                else { rawB[c][0] = v.x; rawB[c][F / 2 - 1] = v.y; }
            } else {
                const uint32_t v = __ldg(reinterpret_cast<const uint32_t*>(table + static_cast<size_t>(idx) * 2));
                if (it == 0) rawA[c][0] = v; else rawB[c][0] = v;
            }
        }
    }
    // exchange: the even lane needs A(dx=1) from the odd lane, the odd lane needs B(dx=0) from the even lane
    uint32_t own[4][F / 2], got[4][F / 2];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int k = 0; k < F / 2; ++k) {
            const uint32_t send = odd ? rawA[c][k] : rawB[c][k];
            own[c][k] = odd ? rawB[c][k] : rawA[c][k];
            got[c][k] = __shfl_xor_sync(full, send, 1);
        }
    __half2 acc[F / 2];
#pragma unroll
    for (int k = 0; k < F / 2; ++k) acc[k] = __float2half2_rn(0.0f);
#pragma unroll
    for (int c = 0; c < 8; ++c) {  // tiny-cuda-nn corner order: bit 0 = x, bit 1 = y, bit 2 = z
        float wt = 1.0f;
#pragma unroll
        for (int d = 0; d < 3; ++d) wt *= (c & (1 << d)) ? w[d] : 1.0f - w[d];
        const __half2 wh = __half2half2(__float2half_rn(wt));
        const bool dx = c & 1;
        const int q = c >> 1;  // (dy, dz) combination
#pragma unroll
        for (int k = 0; k < F / 2; ++k) {
            // own[] holds this lane's dx == odd corners, got[] the dx == !odd corners
            const uint32_t v = (dx == static_cast<bool>(odd)) ? own[q][k] : got[q][k];
            acc[k] = __hfma2(wh, *reinterpret_cast<const __half2*>(&v), acc[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < F / 2; ++k) *reinterpret_cast<__half2*>(dst + 2 * k) = acc[k];
}

// sh_enc (common_device.h:476-502) up to degree 4, fp32 -> fp16.
__device__ __forceinline__ void encode_sh(int degree, const float* d, __half* dst) {
    const float x = d[0] * 2.f - 1.f, y = d[1] * 2.f - 1.f, z = d[2] * 2.f - 1.f;
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    float o[16];
    o[0] = 0.28209479177387814f;
    o[1] = -0.48860251190291987f * y;
    o[2] = 0.48860251190291987f * z;
    o[3] = -0.48860251190291987f * x;
    o[4] = 1.0925484305920792f * xy;
    o[5] = -1.0925484305920792f * yz;
    o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
    o[7] = -1.0925484305920792f * xz;
    o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
    o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
    o[10] = 2.8906114426405538f * xy * z;
    o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
    o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
    o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
    o[14] = 1.4453057213202769f * z * (x2 - y2);
    o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
    const int n = degree * degree;
#pragma unroll
    for (int i = 0; i < 16; ++i)
        if (i < n) dst[i] = __float2half_rn(o[i]);
}

}  // namespace nrb
