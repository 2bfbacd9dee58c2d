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
#include "nrb_frame.cuh"
#include "nrb_query_warp.cuh"

#include <cuda_runtime.h>

namespace nrb {

namespace {

using namespace qwarp;

template <int WIDTH>
__global__ void __launch_bounds__(512, 1)
    k_neural_query(const __grid_constant__ NetDev net, const float* __restrict__ in, int n, const int* dCount,
                   __half* __restrict__ out, int ek, int warps) {
    extern __shared__ __align__(16) unsigned char smem[];
    const int H = net.nHidden, D = net.inputDims;
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
    preload_weights<WIDTH>(net, ek, w0, wh, wl);
    __syncthreads();

    const size_t perWarpBytes = static_cast<size_t>(kRows) * ldA * 2 + static_cast<size_t>(kRows) * 16 * 4;
    __half* act = reinterpret_cast<__half*>(perWarp + warp * perWarpBytes);
    float* inp = reinterpret_cast<float*>(act + kRows * ldA);  // [16][16] floats (D <= 16)
    const int tiles = (n + kRows - 1) / kRows;
    for (int tile = blockIdx.x * warps + warp; tile < tiles; tile += gridDim.x * warps) {
        const int q0 = tile * kRows;
        // stage the tile's inputs (rows past n read as zeros, like the reference's padded tail)
        for (int i = lane; i < kRows * D; i += 32) {
            const int r = i / D, c = i - r * D;
            inp[r * 16 + c] = (q0 + r) < n ? in[static_cast<size_t>(q0 + r) * D + c] : 0.0f;
        }
        __syncwarp();
        query_tile<WIDTH>(net, ek, w0, wh, wl, act, inp, out + static_cast<size_t>(q0) * 16, 16, n - q0);
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

void init_query_kernels() {
    cudaFuncSetAttribute(k_neural_query<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    cudaFuncSetAttribute(k_neural_query<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    init_query_tc_kernels();
    init_query_split_kernels();
}

int launch_neural_query(const NetDev& net, const float* in, size_t n, const int* dCount, __half* out,
                        cudaStream_t s) {
    if (n == 0) return 0;
    const int which = net.kernelChoice & 0xff;
    // automatic choice: the split pays when a lane's encoder task carries enough gathers to amortise its
    // fixed cost (measured: 3 points x F4 -> 2.7x the fused kernel; 1 point x F2 -> 0.5x)
    const bool splitPays = net.nPoints * net.nFeatures >= 8;
    if ((which == 3 || (which == 0 && splitPays)) && net.xtRows > 0 && net.packedW && split_query_supported(net)) {
        NetDev tmp = net;
        tmp.debug = net.kernelChoice >> 8;
        return launch_split_query(tmp, in, n, dCount, out, s);
    }
    // shapes whose tables do not fit shared memory are bound by L2 gathers (BASELINE config 5: 16 levels x 2^19 entries,
    // 512 B gathered against 14-45 KFLOP per query): the warp-autonomous mma.sync kernel keeps many more independent
    // gather streams in flight than the warp-specialised tcgen05 kernel and is the faster of the two there
    // (2^22 queries, width 64 / 128: 2.85 / 2.32 G queries/s against 2.06 / 1.98; tiny-cuda-nn 2.30 / 1.51)
    const bool gatherBound = which == 0 && !splitPays && plan_smem(net).warps > 0;
    if (which != 1 && !gatherBound && neural_query_tc_supported(net)) {
        NetDev tmp = net;
        tmp.debug = net.kernelChoice >> 8;
        return launch_neural_query_tc(tmp, in, n, dCount, out, s);
    }
    const SmemPlan p = plan_smem(net);
    if (p.warps <= 0) return -1;
    const int tiles = static_cast<int>((n + kRows - 1) / kRows);
    int grid = (tiles + p.warps - 1) / p.warps;
    if (grid > sm_count()) grid = sm_count();  // one persistent CTA per SM (weights stay resident in its smem)
    const int threads = p.warps * 32;
    if (net.width == 128) {
        k_neural_query<128><<<grid, threads, p.total, s>>>(net, in, static_cast<int>(n), dCount, out, p.ek, p.warps);
        NRB_LAUNCHED("k_neural_query", s);
    } else {
        k_neural_query<64><<<grid, threads, p.total, s>>>(net, in, static_cast<int>(n), dCount, out, p.ek, p.warps);
        NRB_LAUNCHED("k_neural_query", s);
    }
    return 1;
}

}  // namespace nrb
