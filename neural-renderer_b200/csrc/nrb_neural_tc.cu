// nrb_neural_tc.cu — the neural query on Blackwell's 5th-generation tensor cores (tcgen05 + TMEM).
//
// Same arithmetic contract as nrb_neural.cu (tiny-cuda-nn encodings bit-faithful, fp16 operands,
// fp32 accumulation, fp16 rounding + fp16 LeakyReLU per layer); different machine mapping:
//
//   * tile = 128 queries = the M of one UMMA (cta_group::1, M=128): accumulators are 128 TMEM lanes
//     x N fp32 columns, one query per lane;
//   * all layer weights live in shared memory for the CTA's lifetime as K-major, un-swizzled
//     "core matrix" operands (8 rows x 16 bytes contiguous; LBO = 128 B between the two K halves of
//     one K=16 step, SBO = K*16 B between 8-row groups) and are read by the tensor core directly
//     through the async proxy - no ldmatrix traffic through the LSU, which is what bounded the
//     mma.sync kernel (ncu: L1/TEX 69 % busy, tensor pipe 29 %);
//   * warp-specialised, persistent CTA of 28 warps: warps 4..27 gather and encode tile i+1 straight
//     into the layer-0 A operand X while warps 0..3 run the MLP of tile i: one elected thread issues
//     tcgen05.mma (7 or 8 K-steps per layer) and commits to an mbarrier; the 128 epilogue threads
//     read their query's row from TMEM (tcgen05.ld 32x32b), apply fp16 LeakyReLU and write the next
//     layer's A operand H in place; the last layer (N = 16) goes to global memory as 16 halfs.
//
// mbarrier waits are bounded (trap after ~1 s) so a protocol bug aborts the launch instead of
// hanging the device.
#include <cuda_runtime.h>

#include "nrb_encode.cuh"
#include "nrb_frame.cuh"

namespace nrb {

namespace {

constexpr int kTileM = 128;
constexpr int kMlpThreads = 128;    // warps 0..3: TMEM lane quarter = warp index
constexpr int kEncThreads = 768;    // warps 4..27
constexpr int kInpStride = 17;      // floats per staged input row (odd stride: conflict-free)

__device__ __forceinline__ uint32_t s32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// ---- mbarrier ----------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = s32(bar);
    for (uint32_t spin = 0;; ++spin) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(addr), "r"(parity) : "memory");
        if (done) return;
        if (spin > (1u << 24)) __trap();  // protocol bug: abort the launch, never hang the GPU
    }
}

// ---- tcgen05 -----------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void proxy_fence() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(bar)) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]^T, fp16 operands, fp32 accumulate.
__device__ __forceinline__ void tc_mma_f16(uint32_t dTmem, uint64_t aDesc, uint64_t bDesc, uint32_t iDesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(dTmem), "l"(aDesc), "l"(bDesc), "r"(iDesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Byte offset of element (row, k) of a K-major, un-swizzled UMMA operand with K columns.
__host__ __device__ __forceinline__ uint32_t canon_off(int row, int k, int K) {
    return static_cast<uint32_t>((row >> 3) * (K * 16) + (k >> 3) * 128 + (row & 7) * 16 + (k & 7) * 2);
}
// Shared-memory matrix descriptor (SM100 UMMA): start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46),
// version 1 at [46,48), layout type 0 (no swizzle) at [61,64).
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, int K) {
    return static_cast<uint64_t>((addr & 0x3FFFFu) >> 4) | (static_cast<uint64_t>(128 >> 4) << 16) |
           (static_cast<uint64_t>((K * 16) >> 4) << 32) | (1ull << 46);
}
// Instruction descriptor, kind::f16: D = F32 (bit 4), A = B = F16 (0), both K-major, N>>3 at 17, M>>4 at 24.
__host__ __device__ __forceinline__ uint32_t instr_desc(int N) {
    return (1u << 4) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(kTileM >> 4) << 24);
}

struct TcPlan {
    int ek;
    uint32_t offW0, offWh, offWl, offX, offH, offInp, offBar, total;
};

TcPlan plan_tc(const NetDev& net) {
    TcPlan p{};
    p.ek = (net.encDims + 15) & ~15;
    const uint32_t W = net.width;
    uint32_t o = 0;
    p.offW0 = o; o += W * p.ek * 2;
    p.offWh = o; o += (net.nHidden - 1) * W * W * 2;
    p.offWl = o; o += 16 * W * 2;
    p.offX = o; o += kTileM * p.ek * 2;
    p.offH = o; o += kTileM * W * 2;
    p.offInp = o; o += kTileM * kInpStride * 4;
    p.offBar = o; o += 64;
    p.total = o;
    return p;
}

// Copy one weight matrix W[rows][cols] (row-major, the tiny-cuda-nn parameter layout) into the UMMA
// K-major operand layout with K = kPad columns (zero-filled past cols).
__device__ __forceinline__ void stage_weights(unsigned char* dst, const __half* __restrict__ src, int rows, int cols,
                                              int kPad, int tid, int nthreads) {
    const int chunksPerRow = kPad >> 3;
    for (int i = tid; i < rows * chunksPerRow; i += nthreads) {
        const int r = i / chunksPerRow, c8 = (i - r * chunksPerRow) << 3;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (c8 + 8 <= cols && ((cols & 7) == 0)) {
            v = __ldg(reinterpret_cast<const uint4*>(src + static_cast<size_t>(r) * cols + c8));
        } else {
            __half tmp[8];
            for (int j = 0; j < 8; ++j) tmp[j] = c8 + j < cols ? src[static_cast<size_t>(r) * cols + c8 + j] : __float2half_rn(0.0f);
            v = *reinterpret_cast<const uint4*>(tmp);
        }
        *reinterpret_cast<uint4*>(dst + canon_off(r, c8, kPad)) = v;
    }
}

__device__ __forceinline__ uint32_t pack_leaky(uint32_t a, uint32_t b) {
    return leaky_pack2(__uint_as_float(a), __uint_as_float(b));
}
__device__ __forceinline__ uint32_t pack_plain(uint32_t a, uint32_t b) {
    const __half2 h = __floats2half2_rn(__uint_as_float(a), __uint_as_float(b));
    return *reinterpret_cast<const uint32_t*>(&h);
}

template <int WIDTH>
__global__ void __launch_bounds__(kMlpThreads + kEncThreads, 1)
    k_neural_query_tc(const __grid_constant__ NetDev net, const float* __restrict__ in, int n, const int* dCount,
                      __half* __restrict__ out, const __grid_constant__ TcPlan plan) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (dCount) {
        const int dn = *dCount;
        n = dn < n ? dn : n;
    }
    const int tiles = (n + kTileM - 1) / kTileM;
    if (static_cast<int>(blockIdx.x) >= tiles) return;  // nothing for this CTA: skip preload and TMEM

    const int E = net.encDims, EK = plan.ek, H = net.nHidden, D = net.inputDims;
    unsigned char* w0 = smem + plan.offW0;
    unsigned char* wh = smem + plan.offWh;
    unsigned char* wl = smem + plan.offWl;
    unsigned char* X = smem + plan.offX;
    unsigned char* Hb = smem + plan.offH;
    float* inp = reinterpret_cast<float*>(smem + plan.offInp);
    uint64_t* barXFull = reinterpret_cast<uint64_t*>(smem + plan.offBar);
    uint64_t* barXEmpty = barXFull + 1;
    uint64_t* barMma = barXFull + 2;
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(barXFull + 3);

    // ---- one-time setup ---------------------------------------------------------------------------
    if (tid == 0) {
        mbar_init(barXFull, kEncThreads / 32);
        mbar_init(barXEmpty, 1);
        mbar_init(barMma, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(tmemSlot, 2 * WIDTH);  // two accumulators of WIDTH fp32 columns
    {
        const __half* src = net.mlp;
        stage_weights(w0, src, WIDTH, E, EK, tid, blockDim.x);
        src += WIDTH * E;
        for (int l = 0; l < H - 1; ++l) {
            stage_weights(wh + l * WIDTH * WIDTH * 2, src, WIDTH, WIDTH, WIDTH, tid, blockDim.x);
            src += WIDTH * WIDTH;
        }
        stage_weights(wl, src, 16, WIDTH, WIDTH, tid, blockDim.x);
        // X columns past the encoding (K padding) stay zero for the kernel's lifetime
        for (int i = tid; i < kTileM * EK / 8; i += blockDim.x) reinterpret_cast<uint4*>(X)[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    proxy_fence();       // weights were written through the generic proxy, the tensor core reads them
    tc_fence_before();   // through the async proxy
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmemSlot;

    if (warp >= kMlpThreads / 32) {
        // ===================== encoder warps: gather + encode tile i into X ==========================
        const int et = tid - kMlpThreads;
        const int PL = net.nPoints * net.nLevels, F = net.nFeatures;
        const int shBase = PL * F, shN = net.shDegree * net.shDegree;
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int q0 = tile * kTileM;
            mbar_wait(barXEmpty, (it & 1) ^ 1);  // layer 0 of the previous tile has consumed X
            for (int i = et; i < kTileM * D; i += kEncThreads) {
                const int r = i / D, c = i - r * D;
                inp[r * kInpStride + c] = (q0 + r) < n ? __ldg(in + static_cast<size_t>(q0 + r) * D + c) : 0.0f;
            }
            asm volatile("bar.sync 2, %0;" ::"n"(kEncThreads) : "memory");
            for (int item = et; item < kTileM * PL; item += kEncThreads) {
                const int pl = item >> 7, r = item & (kTileM - 1);  // consecutive lanes: same level, next query
                const int p = pl / net.nLevels, level = pl - p * net.nLevels;
                const float* x = inp + r * kInpStride + 3 * p;
                __half* dst = reinterpret_cast<__half*>(X + canon_off(r, pl * F, EK));
                if (F == 4) encode_item_paired<4>(net, level, x, dst);
                else encode_item_paired<2>(net, level, x, dst);
            }
            if (shN > 0 && et < kTileM) {
                __half sh[16];
                encode_sh(net.shDegree, inp + et * kInpStride + 3 * net.nPoints, sh);
                for (int c = 0; c < shN; ++c) *reinterpret_cast<__half*>(X + canon_off(et, shBase + c, EK)) = sh[c];
            }
            proxy_fence();  // make this thread's X writes visible to the tensor core
            __syncwarp();
            if (lane == 0) mbar_arrive(barXFull);
            asm volatile("bar.sync 2, %0;" ::"n"(kEncThreads) : "memory");  // inp is reused by the next tile
        }
    } else {
        // ===================== MLP warps: tcgen05.mma issue + TMEM epilogues =========================
        const uint32_t laneBase = static_cast<uint32_t>(warp * 32) << 16;  // this warp's TMEM lane quarter
        const int row = tid;                                                 // query row of this thread
        const uint32_t descX = s32(X), descH = s32(Hb);
        uint32_t mmaPhase = 0;
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int q0 = tile * kTileM;
            mbar_wait(barXFull, it & 1);
            tc_fence_after();
            if (net.debug & 2) {
                if (tid == 0) mbar_arrive(barXEmpty);
                continue;
            }
            if (tid == 0) {
                const uint32_t idesc = instr_desc(WIDTH);
                const uint64_t a0 = smem_desc(descX, EK), b0 = smem_desc(s32(w0), EK);
                for (int k = 0; k < EK / 16; ++k) tc_mma_f16(tmem, a0 + 16u * k, b0 + 16u * k, idesc, k > 0);
                tc_commit(barXEmpty);  // X may be refilled as soon as these MMAs have read it
                tc_commit(barMma);
            }
            for (int layer = 0; layer < H; ++layer) {
                mbar_wait(barMma, mmaPhase);
                mmaPhase ^= 1u;
                tc_fence_after();
                const uint32_t acc = tmem + laneBase + static_cast<uint32_t>((layer & 1) * WIDTH);
#pragma unroll 1
                for (int c0 = 0; c0 < WIDTH; c0 += 32) {
                    uint32_t v[32];
                    tc_ld32(acc + c0, v);
                    tc_wait_ld();
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        uint4 o;
                        o.x = pack_leaky(v[8 * j + 0], v[8 * j + 1]);
                        o.y = pack_leaky(v[8 * j + 2], v[8 * j + 3]);
                        o.z = pack_leaky(v[8 * j + 4], v[8 * j + 5]);
                        o.w = pack_leaky(v[8 * j + 6], v[8 * j + 7]);
                        *reinterpret_cast<uint4*>(Hb + canon_off(row, c0 + 8 * j, WIDTH)) = o;
                    }
                }
                proxy_fence();
                tc_fence_before();
                asm volatile("bar.sync 1, %0;" ::"n"(kMlpThreads) : "memory");
                if (tid == 0) {
                    tc_fence_after();
                    const bool last = layer == H - 1;
                    const uint32_t idesc = instr_desc(last ? 16 : WIDTH);
                    const unsigned char* wsm = last ? wl : wh + layer * WIDTH * WIDTH * 2;
                    const uint64_t a0 = smem_desc(descH, WIDTH), b0 = smem_desc(s32(wsm), WIDTH);
                    const uint32_t d = tmem + static_cast<uint32_t>(((layer + 1) & 1) * WIDTH);
                    for (int k = 0; k < WIDTH / 16; ++k) tc_mma_f16(d, a0 + 16u * k, b0 + 16u * k, idesc, k > 0);
                    tc_commit(barMma);
                }
            }
            // output layer: 16 fp32 columns -> 16 halfs per query
            mbar_wait(barMma, mmaPhase);
            mmaPhase ^= 1u;
            tc_fence_after();
            {
                uint32_t v[16];
                tc_ld16(tmem + laneBase + static_cast<uint32_t>((H & 1) * WIDTH), v);
                tc_wait_ld();
                if (q0 + row < n) {
                    uint4 lo, hi;
                    lo.x = pack_plain(v[0], v[1]); lo.y = pack_plain(v[2], v[3]);
                    lo.z = pack_plain(v[4], v[5]); lo.w = pack_plain(v[6], v[7]);
                    hi.x = pack_plain(v[8], v[9]); hi.y = pack_plain(v[10], v[11]);
                    hi.z = pack_plain(v[12], v[13]); hi.w = pack_plain(v[14], v[15]);
                    uint4* dst = reinterpret_cast<uint4*>(out + static_cast<size_t>(q0 + row) * 16);
                    dst[0] = lo;
                    dst[1] = hi;
                }
            }
            tc_fence_before();
            asm volatile("bar.sync 1, %0;" ::"n"(kMlpThreads) : "memory");  // accumulators free for the next tile
        }
    }
    // ---- teardown ---------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 2 * WIDTH);
}

}  // namespace

bool neural_query_tc_supported(const NetDev& net) {
    if (net.width != 64 && net.width != 128) return false;
    if (net.nFeatures != 2 && net.nFeatures != 4) return false;
    if (net.inputDims >= kInpStride || net.nHidden < 1 || net.shDegree > 4) return false;
    if (((net.nPoints * net.nLevels * net.nFeatures) & 7) != 0 && net.shDegree > 0) return false;
    return plan_tc(net).total <= 227u * 1024u;
}

void init_query_tc_kernels() {
    cudaFuncSetAttribute(k_neural_query_tc<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    cudaFuncSetAttribute(k_neural_query_tc<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
}

int launch_neural_query_tc(const NetDev& net, const float* in, size_t n, const int* dCount, __half* out,
                           cudaStream_t s) {
    if (n == 0) return 0;
    const TcPlan p = plan_tc(net);
    const int tiles = static_cast<int>((n + kTileM - 1) / kTileM);
    const int grid = tiles < sm_count() ? tiles : sm_count();  // one persistent CTA per SM
    if (net.width == 128) {
        k_neural_query_tc<128><<<grid, kMlpThreads + kEncThreads, p.total, s>>>(net, in, static_cast<int>(n), dCount, out, p);
    } else {
        k_neural_query_tc<64><<<grid, kMlpThreads + kEncThreads, p.total, s>>>(net, in, static_cast<int>(n), dCount, out, p);
    }
    NRB_LAUNCHED("k_neural_query_tc", s);
    return 1;
}

}  // namespace nrb
