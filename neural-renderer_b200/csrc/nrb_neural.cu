// nrb_neural.cu — the neural query: hash-grid + SH encoding and the fused MLP in ONE kernel.
//
// Reference arithmetic (vendored tiny-cuda-nn, julcst fork @2a4340e under ext/tiny-cuda-nn):
//   kernel_grid            include/tiny-cuda-nn/encodings/grid.h:48-169 (pos_fract, grid_index,
//                          coherent_prime_hash: common_device.h:768-791, 848-895, 1017-1029)
//   kernel_sh / sh_enc     encodings/spherical_harmonics.h:46-72, common_device.h:476-502
//   kernel_mlp_fused       src/fully_fused_mlp.cu:499-557 (weights row-major [out][in], :659-671;
//                          LeakyReLU slope 0.01 applied in fp16, common_device.h:123-129)
// The reference runs 5 launches per query batch (3 x kernel_grid, kernel_sh, kernel_mlp_fused)
// with a [112 x N] fp16 SoA temporary written and re-read through HBM (448 B/query).  Here the
// encoded row goes straight into the shared-memory A-operand of the first layer and never
// leaves the SM.
//
// This file holds the warp-autonomous kernel: each warp owns 16 queries end to end (gathers ->
// A tile in smem -> all layers with mma.sync.m16n8k16 f16 x f16 -> f32 -> 16 outputs), all layer
// weights stay resident in shared memory for the CTA's lifetime, and there is no block-level
// barrier after the weight preload, so gather latency of one warp overlaps tensor work of the
// others.  Accumulation is fp32 (tiny-cuda-nn accumulates in fp16; fp32 is closer to exact).
// The tcgen05/TMEM variant lives in nrb_neural_tc.cu.
#include "nrb_encode.cuh"
#include "nrb_frame.cuh"

#include <cuda_runtime.h>

namespace nrb {

namespace {

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

SmemPlan plan_smem(const NetDev& net) {
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

template <int WIDTH>
__global__ void __launch_bounds__(512, 1)
    k_neural_query(const __grid_constant__ NetDev net, const float* __restrict__ in, int n, const int* dCount,
                   __half* __restrict__ out, int ek, int warps) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int E = net.encDims, H = net.nHidden, D = net.inputDims;
    const int ld0 = ek + 8, ldw = WIDTH + 8, ldA = (ek > WIDTH ? ek : WIDTH) + 8;
    __half* w0 = reinterpret_cast<__half*>(smem);
    __half* wh = w0 + WIDTH * ld0;
    __half* wl = wh + (H - 1) * WIDTH * ldw;
    unsigned char* perWarp = reinterpret_cast<unsigned char*>(wl + 16 * ldw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (dCount) {
        const int dn = *dCount;
        n = dn < n ? dn : n;
    }
    if (blockIdx.x * warps * kRows >= n) return;  // no tile for this CTA: skip the weight preload

    // ---- preload all layer weights (zero-padded k columns) --------------------------------------
    {
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
    __syncthreads();

    const size_t perWarpBytes = static_cast<size_t>(kRows) * ldA * 2 + static_cast<size_t>(kRows) * 16 * 4;
    __half* act = reinterpret_cast<__half*>(perWarp + warp * perWarpBytes);
    float* inp = reinterpret_cast<float*>(act + kRows * ldA);  // [16][16] floats (D <= 16)
    const int PL = net.nPoints * net.nLevels, F = net.nFeatures;
    const int shBase = PL * F, shN = net.shDegree * net.shDegree;
    const int tiles = (n + kRows - 1) / kRows;

    for (int tile = blockIdx.x * warps + warp; tile < tiles; tile += gridDim.x * warps) {
        const int q0 = tile * kRows;
        // stage the tile's inputs (rows past n read as zeros, like the reference's padded tail)
        for (int i = lane; i < kRows * D; i += 32) {
            const int r = i / D, c = i - r * D;
            inp[r * 16 + c] = (q0 + r) < n ? in[static_cast<size_t>(q0 + r) * D + c] : 0.0f;
        }
        __syncwarp();
        // ---- encode: (point, level) major so a half-warp shares one level's table --------------
        for (int it = lane; it < kRows * PL; it += 32) {
            const int pl = it >> 4, r = it & 15;
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
                if (q0 + r0 < n)
                    *reinterpret_cast<__half2*>(out + static_cast<size_t>(q0 + r0) * 16 + j * 8 + c0) =
                        __floats2half2_rn(acc[j][0], acc[j][1]);
                if (q0 + r0 + 8 < n)
                    *reinterpret_cast<__half2*>(out + static_cast<size_t>(q0 + r0 + 8) * 16 + j * 8 + c0) =
                        __floats2half2_rn(acc[j][2], acc[j][3]);
            }
        }
        __syncwarp();
    }
}

}  // namespace

size_t neural_query_smem(const NetDev& net, int* warpsOut) {
    if ((net.width != 64 && net.width != 128) || (net.nFeatures != 2 && net.nFeatures != 4) || net.inputDims > 16 ||
        net.nHidden < 1 || net.shDegree > 4)
        return 0;
    const SmemPlan p = plan_smem(net);
    if (warpsOut) *warpsOut = p.warps;
    return p.warps > 0 ? p.total : 0;
}

static int g_queryKernel = 0;
void set_query_kernel(int which) { g_queryKernel = which; }
int query_kernel() { return g_queryKernel; }

int launch_neural_query(const NetDev& net, const float* in, size_t n, const int* dCount, __half* out,
                        cudaStream_t s) {
    if (n == 0) return 0;
    const int which = g_queryKernel & 0xff;
    // automatic choice: the split pays when a lane's encoder task carries enough gathers to amortise its
    // fixed cost (measured: 3 points x F4 -> 2.7x the fused kernel; 1 point x F2 -> 0.5x)
    const bool splitPays = net.nPoints * net.nFeatures >= 8;
    if ((which == 3 || (which == 0 && splitPays)) && net.xtRows > 0 && net.packedW && split_query_supported(net)) {
        NetDev tmp = net;
        tmp.debug = g_queryKernel >> 8;
        return launch_split_query(tmp, in, n, dCount, out, s);
    }
    if (which != 1 && neural_query_tc_supported(net)) {
        NetDev tmp = net;
        tmp.debug = g_queryKernel >> 8;
        return launch_neural_query_tc(tmp, in, n, dCount, out, s);
    }
    const SmemPlan p = plan_smem(net);
    if (p.warps <= 0) return -1;
    const int tiles = static_cast<int>((n + kRows - 1) / kRows);
    int grid = (tiles + p.warps - 1) / p.warps;
    if (grid > 148) grid = 148;  // one persistent CTA per SM (weights stay resident in its smem)
    const int threads = p.warps * 32;
    if (net.width == 128) {
        static bool attr = false;
        if (!attr) {
            cudaFuncSetAttribute(k_neural_query<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
            attr = true;
        }
        k_neural_query<128><<<grid, threads, p.total, s>>>(net, in, static_cast<int>(n), dCount, out, p.ek, p.warps);
        NRB_LAUNCHED("k_neural_query", s);
    } else {
        static bool attr = false;
        if (!attr) {
            cudaFuncSetAttribute(k_neural_query<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
            attr = true;
        }
        k_neural_query<64><<<grid, threads, p.total, s>>>(net, in, static_cast<int>(n), dCount, out, p.ek, p.warps);
        NRB_LAUNCHED("k_neural_query", s);
    }
    return 1;
}

}  // namespace nrb
