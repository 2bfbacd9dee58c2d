// nrb_query_warp.cuh — the warp-autonomous neural query as device functions: one warp evaluates the network
// for a tile of 16 queries (gathers -> A tile in shared memory -> every layer with mma.sync.m16n8k16
// f16 x f16 -> f32 -> 16 outputs) against layer weights that its CTA keeps resident in shared memory.
// Used by the one-launch query kernel (nrb_neural.cu) and by the fused segment-loop tail
// (nrb_segments.cu: k_seg_tail), where a warp carries 16 rays through exits -> query -> apply per iteration.
#pragma once

#include "nrb_encode.cuh"

namespace nrb {
namespace qwarp {

constexpr int kRows = 16;  // queries per warp tile (one m16 MMA row block)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}

__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

struct SmemPlan {
    int ek, ld0, ldw, ldA, warps;
    size_t weightHalfs, perWarpBytes, total;
};

inline SmemPlan plan_smem(const NetDev& net) {
    SmemPlan p{};
    p.ek = (net.encDims + 15) & ~15;
    p.ld0 = p.ek + 8;
    p.ldw = net.width + 8;
    p.ldA = (p.ek > net.width ? p.ek : net.width) + 8;
    p.weightHalfs = static_cast<size_t>(net.width) * p.ld0 + static_cast<size_t>(net.nHidden - 1) * net.width * p.ldw +
                    16u * p.ldw;
    p.perWarpBytes = static_cast<size_t>(kRows) * p.ldA * 2 + static_cast<size_t>(kRows) * 16 * 4;  // act + inputs
    const size_t budget = 227 * 1024;
    const size_t wbytes = p.weightHalfs * 2;
    int warps = wbytes + p.perWarpBytes > budget ? 0 : static_cast<int>((budget - wbytes) / p.perWarpBytes);
    if (warps > 16) warps = 16;
    p.warps = warps;
    p.total = wbytes + p.perWarpBytes * (warps > 0 ? warps : 0);
    return p;
}

// Copy all layer weights into the CTA's shared memory (zero-padded k columns); every thread of the CTA.
template <int WIDTH>
__device__ __forceinline__ void preload_weights(const NetDev& net, int ek, __half* w0, __half* wh, __half* wl) {
    const int E = net.encDims, H = net.nHidden;
    const int ld0 = ek + 8, ldw = WIDTH + 8;
    const __half* src = net.mlp;
    const __half zero = __float2half_rn(0.0f);
    for (int i = threadIdx.x; i < WIDTH * ld0; i += blockDim.x) {
        const int r = i / ld0, c = i - r * ld0;
        w0[i] = c < E ? src[r * E + c] : zero;
    }
    src += WIDTH * E;
    for (int l = 0; l < H - 1; ++l) {
        for (int i = threadIdx.x; i < WIDTH * ldw; i += blockDim.x) {
            const int r = i / ldw, c = i - r * ldw;
            wh[l * WIDTH * ldw + i] = c < WIDTH ? src[r * WIDTH + c] : zero;
        }
        src += WIDTH * WIDTH;
    }
    for (int i = threadIdx.x; i < 16 * ldw; i += blockDim.x) {
        const int r = i / ldw, c = i - r * ldw;
        wl[i] = c < WIDTH ? src[r * WIDTH + c] : zero;
    }
}

// One tile: inp [16][16] fp32 inputs (staged by the caller, rows past the batch zero) -> out rows of 16 halfs
// at out + row * outStride for rows < validRows.  act is the warp's [16][ldA] fp16 scratch.  Whole warp.
template <int WIDTH>
__device__ __forceinline__ void query_tile(const NetDev& net, int ek, const __half* w0, const __half* wh, const __half* wl,
                                           __half* act, const float* inp, __half* out, size_t outStride, int validRows) {
    const int E = net.encDims, H = net.nHidden;
    const int ld0 = ek + 8, ldw = WIDTH + 8, ldA = (ek > WIDTH ? ek : WIDTH) + 8;
    const int lane = threadIdx.x & 31;
    const int PL = net.nPoints * net.nLevels, F = net.nFeatures;
    const int shBase = PL * F, shN = net.shDegree * net.shDegree;
    // ---- encode: (point, level) major so a half-warp shares one level's table --------------
    // a lane keeps its row for the whole loop (32 is a multiple of 16): its point's coordinates are read when the point
    // changes, not per level - the 16 rows of the [16][16] input tile sit on two banks, an 8-way conflict per read
    int curP = -1;
    float x0 = 0.0f, x1 = 0.0f, x2 = 0.0f;
    for (int it = lane; it < kRows * PL; it += 32) {
        const int pl = it >> 4, r = it & 15;
        const int p = pl / net.nLevels, level = pl - p * net.nLevels;
        if (p != curP) {
            const float* x = inp + r * 16 + 3 * p;
            x0 = x[0]; x1 = x[1]; x2 = x[2];
            curP = p;
        }
        __half* dst = act + r * ldA + pl * F;
        if (F == 4) encode_item<4>(net, level, x0, x1, x2, dst);
        else encode_item<2>(net, level, x0, x1, x2, dst);
    }
    if (lane < kRows) {
        __half* dst = act + lane * ldA + shBase;
        if (shN > 0) encode_sh(net.shDegree, inp + lane * 16 + 3 * net.nPoints, dst);
        for (int c = E; c < ek; ++c) act[lane * ldA + c] = __float2half_rn(0.0f);
    }
    __syncwarp();

    // ---- layers ---------------------------------------------------------------------------
    const uint32_t aBase = smem_u32(act) + ((lane & 15) * ldA + (lane >> 4) * 8) * 2;
    const int bRow = (lane & 7) + (lane >> 4) * 8, bCol = ((lane >> 3) & 1) * 8;
    const __half* W = w0;
    int K = ek, ldW = ld0;
    for (int layer = 0; layer < H; ++layer) {
        float acc[WIDTH / 8][4];
#pragma unroll
        for (int j = 0; j < WIDTH / 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f;
        const uint32_t wBase = smem_u32(W) + (bRow * ldW + bCol) * 2;
        for (int k0 = 0; k0 < K; k0 += 16) {
            uint32_t a0, a1, a2, a3;
            ldsm_x4(aBase + k0 * 2, a0, a1, a2, a3);
#pragma unroll
            for (int nb = 0; nb < WIDTH / 16; ++nb) {
                uint32_t b0, b1, b2, b3;
                ldsm_x4(wBase + (nb * 16 * ldW + k0) * 2, b0, b1, b2, b3);
                mma16816(acc[2 * nb], a0, a1, a2, a3, b0, b1);
                mma16816(acc[2 * nb + 1], a0, a1, a2, a3, b2, b3);
            }
        }
        __syncwarp();  // all A reads of this layer are done before the tile is overwritten
        const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
        for (int j = 0; j < WIDTH / 8; ++j) {
            *reinterpret_cast<uint32_t*>(act + r0 * ldA + j * 8 + c0) = leaky_pack2(acc[j][0], acc[j][1]);
            *reinterpret_cast<uint32_t*>(act + (r0 + 8) * ldA + j * 8 + c0) = leaky_pack2(acc[j][2], acc[j][3]);
        }
        __syncwarp();
        W = layer == 0 ? wh : W + WIDTH * ldw;
        K = WIDTH;
        ldW = ldw;
    }
    // ---- output layer (16 x WIDTH, no activation) ----------------------------------------
    {
        float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
        const uint32_t wBase = smem_u32(wl) + (bRow * ldw + bCol) * 2;
        for (int k0 = 0; k0 < WIDTH; k0 += 16) {
            uint32_t a0, a1, a2, a3, b0, b1, b2, b3;
            ldsm_x4(aBase + k0 * 2, a0, a1, a2, a3);
            ldsm_x4(wBase + k0 * 2, b0, b1, b2, b3);
            mma16816(acc[0], a0, a1, a2, a3, b0, b1);
            mma16816(acc[1], a0, a1, a2, a3, b2, b3);
        }
        const int r0 = lane >> 2, c0 = (lane & 3) * 2;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            if (r0 < validRows)
                *reinterpret_cast<__half2*>(out + static_cast<size_t>(r0) * outStride + j * 8 + c0) = __floats2half2_rn(acc[j][0], acc[j][1]);
            if (r0 + 8 < validRows)
                *reinterpret_cast<__half2*>(out + static_cast<size_t>(r0 + 8) * outStride + j * 8 + c0) = __floats2half2_rn(acc[j][2], acc[j][3]);
        }
    }
    __syncwarp();
}

// The same tile with the layer weights read straight from global memory (row-major [out][in] fp16, the checkpoint's
// layout) instead of a shared-memory copy: a thread's B fragment of mma.m16n8k16 (row.col) is two 32-bit words of
// one weight row, W[n0 + lane / 4][k0 + 2 * (lane % 4) + {0, 1}] and the same at k + 8.  For kernels in which a few
// warps evaluate a few tiles (k_path_tail): no 170 KB preload per CTA, and the shared-memory carve-out stays small,
// so the L1 keeps the BVH nodes the same warps traverse.  Same products, same fp32 summation order as query_tile.
template <int WIDTH>
__device__ __forceinline__ void query_tile_gmem(const NetDev& net, int ek, __half* act, const float* inp, __half* out,
                                                size_t outStride, int validRows, int encodeRows = kRows) {
    const int E = net.encDims, H = net.nHidden;
    const int ldA = (ek > WIDTH ? ek : WIDTH) + 8;
    const int lane = threadIdx.x & 31;
    const int PL = net.nPoints * net.nLevels, F = net.nFeatures;
    const int shBase = PL * F, shN = net.shDegree * net.shDegree;
    // only the first encodeRows rows carry queries (a warp that holds one or two paths): the other rows of the tile
    // are not encoded - rows are independent in the products, whatever they hold never reaches a valid output
    for (int it = lane; it < encodeRows * PL; it += 32) {
        const int pl = it / encodeRows, r = it - pl * encodeRows;
        const int p = pl / net.nLevels, level = pl - p * net.nLevels;
        const float* x = inp + r * 16 + 3 * p;
        __half* dst = act + r * ldA + pl * F;
        if (F == 4) encode_item<4>(net, level, x, dst);
        else encode_item<2>(net, level, x, dst);
    }
    if (lane < kRows) {
        __half* dst = act + lane * ldA + shBase;
        if (shN > 0) encode_sh(net.shDegree, inp + lane * 16 + 3 * net.nPoints, dst);
        for (int c = E; c < ek; ++c) act[lane * ldA + c] = __float2half_rn(0.0f);
    }
    __syncwarp();
    const uint32_t aBase = smem_u32(act) + ((lane & 15) * ldA + (lane >> 4) * 8) * 2;
    const int bn = lane >> 2, bk = (lane & 3) * 2;
    const __half* W = net.mlp;
    int K = ek, Kreal = E;  // layer 0: K padded to 16 in the tile, the weight rows hold E columns
    for (int layer = 0; layer < H; ++layer) {
        float acc[WIDTH / 8][4];
#pragma unroll
        for (int j = 0; j < WIDTH / 8; ++j) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f;
        for (int k0 = 0; k0 < K; k0 += 16) {
            uint32_t a0, a1, a2, a3;
            ldsm_x4(aBase + k0 * 2, a0, a1, a2, a3);
            const int c0 = k0 + bk, c1 = k0 + 8 + bk;
#pragma unroll
            for (int j = 0; j < WIDTH / 8; ++j) {
                const __half* row = W + static_cast<size_t>(j * 8 + bn) * Kreal;
                const uint32_t b0 = c0 < Kreal ? __ldg(reinterpret_cast<const uint32_t*>(row + c0)) : 0u;
                const uint32_t b1 = c1 < Kreal ? __ldg(reinterpret_cast<const uint32_t*>(row + c1)) : 0u;
                mma16816(acc[j], a0, a1, a2, a3, b0, b1);
            }
        }
        __syncwarp();
        const int r0 = lane >> 2, cc = (lane & 3) * 2;
#pragma unroll
        for (int j = 0; j < WIDTH / 8; ++j) {
            *reinterpret_cast<uint32_t*>(act + r0 * ldA + j * 8 + cc) = leaky_pack2(acc[j][0], acc[j][1]);
            *reinterpret_cast<uint32_t*>(act + (r0 + 8) * ldA + j * 8 + cc) = leaky_pack2(acc[j][2], acc[j][3]);
        }
        __syncwarp();
        W += static_cast<size_t>(WIDTH) * Kreal;
        K = WIDTH;
        Kreal = WIDTH;
    }
    {
        float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
        for (int k0 = 0; k0 < WIDTH; k0 += 16) {
            uint32_t a0, a1, a2, a3;
            ldsm_x4(aBase + k0 * 2, a0, a1, a2, a3);
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const __half* row = W + static_cast<size_t>(j * 8 + bn) * WIDTH;
                const uint32_t b0 = __ldg(reinterpret_cast<const uint32_t*>(row + k0 + bk));
                const uint32_t b1 = __ldg(reinterpret_cast<const uint32_t*>(row + k0 + 8 + bk));
                mma16816(acc[j], a0, a1, a2, a3, b0, b1);
            }
        }
        const int r0 = lane >> 2, cc = (lane & 3) * 2;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            if (r0 < validRows)
                *reinterpret_cast<__half2*>(out + static_cast<size_t>(r0) * outStride + j * 8 + cc) = __floats2half2_rn(acc[j][0], acc[j][1]);
            if (r0 + 8 < validRows)
                *reinterpret_cast<__half2*>(out + static_cast<size_t>(r0 + 8) * outStride + j * 8 + cc) = __floats2half2_rn(acc[j][2], acc[j][3]);
        }
    }
    __syncwarp();
}


}  // namespace qwarp
}  // namespace nrb
