// nrb_query_split.cu — the neural query as two kernels sized for the two different machines it needs.
//
// Why split (ncu on the fused kernel, profiles/r1_v2_*): with weights (128 KB) and activation tiles in
// shared memory there is no room left for the hash tables, so every one of the 192 corner fetches per
// query is an L2 sector request (fused kernel: 1.05 G queries/s; 2.5 G/s with the table loads removed).
// The hash grid of the reference network is 8 levels x 128 KB: ONE level fits one SM's shared memory.
//
//   k_query_encode  grid = L x S CTAs, CTA (level, slice): stages its level's table in shared memory
//                   once, then encodes that level for every query of its slice of 128-query tiles
//                   (P points x 8 corners, tiny-cuda-nn arithmetic bit-faithful: fp32 position, u32
//                   coherent-prime hash, fp16 __hfma2 corner chain in corner order) and writes the
//                   level's columns of the encoded tile.  The spherical harmonics and the K padding
//                   are spread over the level CTAs by row (row % L == level).
//   k_query_mlp     persistent, one CTA per SM, all layer weights resident in shared memory as K-major
//                   UMMA operands (pre-packed once at weight load: one bulk copy per CTA).  A producer
//                   thread streams encoded tiles in with cp.async.bulk; one thread issues tcgen05.mma
//                   for TWO tiles in flight, offset by half a tile, so that the TMEM -> LeakyReLU ->
//                   fp16 -> shared-memory epilogue of one tile (4 warps per tile) runs under the MMAs of
//                   the other; accumulators live in TMEM (one WIDTH-column block per tile in flight).
//
// Encoded tile layout (global and shared alike): [tile][k/8][row 0..127][8 halfs] — every 16-byte
// k-chunk of 128 rows is contiguous (2 KB), which is a legal un-swizzled K-major UMMA operand with
// LBO = 2048 (between k-chunks) and SBO = 128 (between 8-row groups), is written by the encoder with
// full-sector stores and is fetched by the MLP kernel as ONE contiguous bulk copy.
// The K order of layer 0 is permuted to level-major (k' = level*(P*F) + point*F + f; tiny-cuda-nn has
// k = (point*L + level)*F + f) so that a level CTA owns a contiguous K range; W0's columns are
// permuted identically at pre-pack time, so the products summed per neuron are the same set.
#include <cuda_runtime.h>

#include <type_traits>

#include "nrb_encode.cuh"
#include "nrb_frame.cuh"
#include "nrb_tc.cuh"

namespace nrb {

namespace {

constexpr int kTileM = 128;
constexpr int kEncThreads = 1024;
constexpr int kEpiWarps = 8;      // per tile in flight: two warps per TMEM lane quarter, half the columns each
constexpr int kMlpThreads = 64 + 2 * kEpiWarps * 32;  // warp 0 producer, warp 1 MMA issuer, then the epilogues of tile A, B
constexpr uint32_t kSmemMax = 227u * 1024u;

// Byte offset of element (row, k) of a weight matrix with K columns in the packed (UMMA K-major,
// un-swizzled, LBO = 128, SBO = K*16) layout.
__host__ __device__ __forceinline__ uint32_t canon_off(int row, int k, int K) {
    return static_cast<uint32_t>((row >> 3) * (K * 16) + (k >> 3) * 128 + (row & 7) * 16 + (k & 7) * 2);
}

struct SplitPlan {
    int ek;                       // layer-0 K, padded to 16
    uint32_t w0Bytes, whBytes, wlBytes, packedBytes;
    uint32_t xBytes, hBytes;      // one encoded tile / one hidden tile in shared memory
    uint32_t offW, offX, offH, offBar, mlpSmem;
    uint32_t encSmem;             // table bytes staged by an encoder CTA (largest level that fits), 128-aligned
    uint32_t encRing;             // per-warp two-stage input rings
    int slices;                   // encoder CTAs per level
};

SplitPlan make_plan(const NetDev& net) {
    SplitPlan p{};
    const uint32_t W = net.width;
    p.ek = (net.encDims + 15) & ~15;
    p.w0Bytes = W * p.ek * 2;
    p.whBytes = (net.nHidden - 1) * W * W * 2;
    p.wlBytes = 16 * W * 2;
    p.packedBytes = p.w0Bytes + p.whBytes + p.wlBytes;
    p.xBytes = kTileM * p.ek * 2;
    p.hBytes = kTileM * W * 2;
    uint32_t o = 0;
    p.offW = o; o += p.packedBytes;
    p.offX = o; o += p.xBytes;
    p.offH = o; o += 2 * p.hBytes;
    p.offBar = o; o += 128;
    p.mlpSmem = o;
    p.encRing = (kEncThreads / 32) * 2u * 32u * net.inputDims * 4u;
    uint32_t best = 0;
    for (int l = 0; l < net.nLevels; ++l) {
        const uint64_t bytes = (static_cast<uint64_t>(net.levelSize[l]) * net.nFeatures * 2 + 127u) / 128u * 128u;
        if (bytes + p.encRing + 16u <= kSmemMax && bytes > best) best = static_cast<uint32_t>(bytes);
    }
    p.encSmem = best;
    p.slices = net.nLevels > 0 ? (sm_count() / net.nLevels > 0 ? sm_count() / net.nLevels : 1) : 1;
    return p;
}

// ---- weight pre-pack -----------------------------------------------------------------------------
// tiny-cuda-nn parameter layout (row-major [out][in], W0 | W1.. | Wlast) -> the shared-memory image of
// k_query_mlp.  One thread per (matrix row, 8-column chunk).
__global__ void k_pack_weights(const __grid_constant__ NetDev net, unsigned char* dst, int ek) {
    const int W = net.width, E = net.encDims, H = net.nHidden;
    const int P = net.nPoints, L = net.nLevels, F = net.nFeatures, PLF = P * L * F;
    const int c0 = W * (ek >> 3), ch = (H - 1) * W * (W >> 3), cl = 16 * (W >> 3);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < c0 + ch + cl; i += gridDim.x * blockDim.x) {
        __half v[8];
        uint32_t off;
        if (i < c0) {
            const int r = i / (ek >> 3), k8 = (i - r * (ek >> 3)) << 3;
            for (int j = 0; j < 8; ++j) {
                const int kp = k8 + j;  // our K index -> tiny-cuda-nn column
                int c = kp;
                if (kp < PLF) {
                    const int level = kp / (P * F), rem = kp - level * (P * F), p = rem / F, f = rem - p * F;
                    c = (p * L + level) * F + f;
                }
                v[j] = kp < E ? net.mlp[static_cast<size_t>(r) * E + c] : __float2half_rn(0.0f);
            }
            off = canon_off(r, k8, ek);
        } else if (i < c0 + ch) {
            const int t = i - c0, per = W * (W >> 3), l = t / per, u = t - l * per;
            const int r = u / (W >> 3), k8 = (u - r * (W >> 3)) << 3;
            const __half* src = net.mlp + static_cast<size_t>(W) * E + static_cast<size_t>(l) * W * W;
            for (int j = 0; j < 8; ++j) v[j] = src[static_cast<size_t>(r) * W + k8 + j];
            off = static_cast<uint32_t>(W * ek * 2 + l * W * W * 2) + canon_off(r, k8, W);
        } else {
            const int u = i - c0 - ch, r = u / (W >> 3), k8 = (u - r * (W >> 3)) << 3;
            const __half* src = net.mlp + static_cast<size_t>(W) * E + static_cast<size_t>(H - 1) * W * W;
            for (int j = 0; j < 8; ++j) v[j] = src[static_cast<size_t>(r) * W + k8 + j];
            off = static_cast<uint32_t>(W * ek * 2 + (H - 1) * W * W * 2) + canon_off(r, k8, W);
        }
        *reinterpret_cast<uint4*>(dst + off) = *reinterpret_cast<const uint4*>(v);
    }
}

// ---- encoder -------------------------------------------------------------------------------------
template <int F, bool SMEM>
__device__ __forceinline__ void load_entry(const unsigned char* gtable, uint32_t stable, uint32_t byteOff, uint32_t (&raw)[F / 2]) {
    if (SMEM) {
        if (F == 4) {
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(raw[0]), "=r"(raw[F / 2 - 1]) : "r"(stable + byteOff));
        } else {
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(raw[0]) : "r"(stable + byteOff));
        }
    } else {
        if (F == 4) {
            const uint2 v = __ldg(reinterpret_cast<const uint2*>(gtable + byteOff));
            raw[0] = v.x;
            raw[F / 2 - 1] = v.y;
        } else {
            raw[0] = __ldg(reinterpret_cast<const uint32_t*>(gtable + byteOff));
        }
    }
}

struct LevelInfo {
    float scale;
    uint32_t res, size, mask;  // mask = (size - 1) * entryBytes when size is a power of two
    bool dense, pow2;
};

// One (query, point) of this CTA's level: 8 corners, tiny-cuda-nn's arithmetic (see nrb_encode.cuh
// encode_item, to which this is bit-identical; only the address generation is arranged differently).
template <int F, bool SMEM>
__device__ __forceinline__ void encode_point(const LevelInfo& lv, const unsigned char* gtable, uint32_t stable, float x0,
                                             float x1, float x2, uint32_t (&acc)[F / 2]) {
    constexpr uint32_t kShift = F == 4 ? 3 : 2;  // log2(entry bytes)
    float w[3];
    uint32_t g[3];
    const float xs[3] = {x0, x1, x2};
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float p = fmaf(lv.scale, xs[d], 0.5f);
        const float fl = floorf(p);
        g[d] = static_cast<uint32_t>(static_cast<int>(fl));
        w[d] = p - fl;
    }
    uint32_t off[8];
    if (!lv.dense && lv.pow2) {
        // (x ^ y*P1 ^ z*P2) & (size-1), scaled to bytes: mask and scale the three terms once, then one
        // three-input XOR per corner.
        const uint32_t ax0 = (g[0] << kShift) & lv.mask, ax1 = ((g[0] + 1u) << kShift) & lv.mask;
        const uint32_t hy = g[1] * 2654435761u, hz = g[2] * 805459861u;
        const uint32_t ay0 = (hy << kShift) & lv.mask, ay1 = ((hy + 2654435761u) << kShift) & lv.mask;
        const uint32_t az0 = (hz << kShift) & lv.mask, az1 = ((hz + 805459861u) << kShift) & lv.mask;
#pragma unroll
        for (int c = 0; c < 8; ++c) off[c] = ((c & 1) ? ax1 : ax0) ^ ((c & 2) ? ay1 : ay0) ^ ((c & 4) ? az1 : az0);
    } else if (lv.dense && lv.pow2) {
        // (x + y*res + z*res^2) & (size-1), scaled to bytes: shift the three terms once, then an add3 and
        // a mask per corner (level CTAs of a dense level would otherwise be the kernel's stragglers)
        const uint32_t sy = lv.res << kShift, sz = (lv.res * lv.res) << kShift;
        const uint32_t ax0 = g[0] << kShift, ax1 = ax0 + (1u << kShift);
        const uint32_t ay0 = g[1] * sy, ay1 = ay0 + sy;
        const uint32_t az0 = g[2] * sz, az1 = az0 + sz;
#pragma unroll
        for (int c = 0; c < 8; ++c) off[c] = (((c & 1) ? ax1 : ax0) + ((c & 2) ? ay1 : ay0) + ((c & 4) ? az1 : az0)) & lv.mask;
    } else {
#pragma unroll
        for (int c = 0; c < 8; ++c)
            off[c] = level_index(lv.dense, lv.pow2, lv.size, lv.res, g[0] + (c & 1), g[1] + ((c >> 1) & 1), g[2] + ((c >> 2) & 1))
                     << kShift;
    }
    uint32_t raw[8][F / 2];
#pragma unroll
    for (int c = 0; c < 8; ++c) load_entry<F, SMEM>(gtable, stable, off[c], raw[c]);
    __half2 a[F / 2];
#pragma unroll
    for (int k = 0; k < F / 2; ++k) a[k] = __float2half2_rn(0.0f);
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        float wt = 1.0f;
#pragma unroll
        for (int d = 0; d < 3; ++d) wt *= (c & (1 << d)) ? w[d] : 1.0f - w[d];
        const __half2 wh = __half2half2(__float2half_rn(wt));
#pragma unroll
        for (int k = 0; k < F / 2; ++k) a[k] = __hfma2(wh, *reinterpret_cast<const __half2*>(&raw[c][k]), a[k]);
    }
#pragma unroll
    for (int k = 0; k < F / 2; ++k) acc[k] = *reinterpret_cast<const uint32_t*>(&a[k]);
}

// Warp task = (tile, 32-row group): lane = one query.  The lane fetches its whole input row once
// (vector loads), encodes this CTA's level for all P points, and writes the level's P*F columns as
// 16-byte / 8-byte pieces of the tile's k-chunks.  One task in L (warp-uniform) also produces the
// spherical harmonics and the K padding of its rows.
template <int F, bool SMEM>
__device__ __forceinline__ void encode_tasks(const NetDev& net, const LevelInfo& lv, int level, const unsigned char* gtable,
                                             uint32_t stable, const float* __restrict__ in, int rowBase, int rowEnd,
                                             int tiles, int* cursor, unsigned char* __restrict__ xt, int ek,
                                             float* stageBase) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int P = net.nPoints, D = net.inputDims, L = net.nLevels;
    // Tasks (tile * 4 + 32-row group) of this level are claimed from its cursor, a whole tile at a time in long
    // launches and one group at a time in short ones, where the four groups of a tile should go to four warps.
    const int nTasks = tiles * 4;
    const int K = nTasks >= 8 * (static_cast<int>(gridDim.x) * (static_cast<int>(blockDim.x) >> 5) / L) ? 4 : 1;
    int endTask = 0;
    auto claim = [&]() {  // warp-uniform; -1 when the level is finished
        int t = 0;
        if (lane == 0) t = atomicAdd(cursor, K);
        t = __shfl_sync(0xffffffffu, t, 0);
        endTask = t + K < nTasks ? t + K : nTasks;
        return t < nTasks ? t : -1;
    };
    const int PF = P * F, PLF = PF * L, shN = net.shDegree * net.shDegree;
    const uint32_t tileBytes = static_cast<uint32_t>(ek) * kTileM * 2;
    const int k0 = level * PF;  // first column of this level
    const bool lPow2 = (L & (L - 1)) == 0;
    auto shOwner = [&](int group) { return (lPow2 ? (group & (L - 1)) : (group % L)) == level; };
    // Input rows reach the lanes through a per-warp two-stage cp.async ring: the 32 rows of a task are
    // one contiguous, 16-byte-aligned block of 128*D bytes, fetched while the previous task computes.
    const int stageFloats = 32 * D;
    float* stage = stageBase + warp * 2 * stageFloats;
    const size_t inBytes = static_cast<size_t>(rowEnd) * D * 4;
    auto fetch = [&](int tile, int group, int slot) {
        const size_t off = static_cast<size_t>(rowBase + tile * kTileM + (group << 5)) * D * 4;
        const unsigned char* src = reinterpret_cast<const unsigned char*>(in) + off;
        const uint32_t dst = tc::s32(stage + slot * stageFloats);
        if (off + 128u * D <= inBytes) {  // the whole 32-row block is inside the batch (all but the last task)
            for (int i = lane; i < 8 * D; i += 32)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * i), "l"(src + 16u * i) : "memory");
        } else {
            for (int i = lane; i < 8 * D; i += 32) {
                const size_t o = off + 16u * i;
                for (int j = 0; j < 4; ++j)
                    stage[slot * stageFloats + 4 * i + j] = o + 4u * j + 4 <= inBytes ? __ldg(in + (o >> 2) + j) : 0.0f;
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int task = claim(), it = 0;
    if (task >= 0) fetch(task >> 2, task & 3, 0);
    for (; task >= 0; ++it) {
        const int curTile = task >> 2, curG = task & 3;
        // the task after this one: the next of the claim, then a new claim
        if (++task >= endTask) task = claim();
        const int slot = it & 1;
        if (task >= 0) {
            fetch(task >> 2, task & 3, slot ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        const int r = (curG << 5) + lane;
        const int row = rowBase + curTile * kTileM + r;
        __syncwarp();  // this task's rows have landed in the slot
        unsigned char* tbase = xt + static_cast<size_t>(curTile) * tileBytes + r * 16;
        float x[16];
        const float* xs = stage + slot * stageFloats + lane * D;
        if ((D & 3) == 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (4 * i < D) {
                    const float4 v = *reinterpret_cast<const float4*>(xs + 4 * i);
                    x[4 * i] = v.x; x[4 * i + 1] = v.y; x[4 * i + 2] = v.z; x[4 * i + 3] = v.w;
                }
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i)
                if (i < D) x[i] = xs[i];
        }
        __syncwarp();  // every lane holds its row: the slot may be refilled by the next fetch
        if (row >= rowEnd) continue;
        uint32_t w[4][F / 2];
#pragma unroll
        for (int p = 0; p < 4; ++p)
            if (p < P) encode_point<F, SMEM>(lv, gtable, stable, x[3 * p], x[3 * p + 1], x[3 * p + 2], w[p]);
        if (F == 4 && P == 3 && (k0 & 3) == 0) {
            // 24 bytes per row: one full 16-byte chunk and half of a neighbouring one
            unsigned char* c0 = tbase + (k0 >> 3) * (kTileM * 16);
            if ((k0 & 7) == 0) {
                *reinterpret_cast<uint4*>(c0) = make_uint4(w[0][0], w[0][F / 2 - 1], w[1][0], w[1][F / 2 - 1]);
                *reinterpret_cast<uint2*>(c0 + kTileM * 16) = make_uint2(w[2][0], w[2][F / 2 - 1]);
            } else {
                *reinterpret_cast<uint2*>(c0 + 8) = make_uint2(w[0][0], w[0][F / 2 - 1]);
                *reinterpret_cast<uint4*>(c0 + kTileM * 16) = make_uint4(w[1][0], w[1][F / 2 - 1], w[2][0], w[2][F / 2 - 1]);
            }
        } else {
#pragma unroll
            for (int p = 0; p < 4; ++p)
                if (p < P) {
                    const int kp = k0 + p * F;
                    unsigned char* dst = tbase + (kp >> 3) * (kTileM * 16) + (kp & 7) * 2;
                    if (F == 4) *reinterpret_cast<uint2*>(dst) = make_uint2(w[p][0], w[p][F / 2 - 1]);
                    else *reinterpret_cast<uint32_t*>(dst) = w[p][0];
                }
        }
        if (shOwner(curTile * 4 + curG) && PLF < ek) {
            // spherical harmonics + K padding: columns [PLF, ek), PLF % 8 == 0 is a support condition
            __half tail[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) tail[i] = __float2half_rn(0.0f);
            if (shN > 0) {
                float d3[3] = {0.0f, 0.0f, 0.0f};
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if (i >= 3 * P && i < 3 * P + 3) d3[i - 3 * P] = x[i];
                encode_sh(net.shDegree, d3, tail);
            }
            for (int k = PLF; k < ek; k += 8)
                *reinterpret_cast<uint4*>(tbase + (k >> 3) * (kTileM * 16)) = *reinterpret_cast<const uint4*>(tail + (k - PLF));
        }
    }
}

template <int F>
__global__ void __launch_bounds__(kEncThreads, 1)
    k_query_encode(const __grid_constant__ NetDev net, const float* __restrict__ in, int n, const int* dCount, int rowBase,
                   int chunkRows, unsigned char* __restrict__ xt, int ek, uint32_t smemBytes) {
    extern __shared__ __align__(128) unsigned char smem[];
    if (dCount) {
        const int dn = *dCount;
        n = dn < n ? dn : n;
    }
    int rowEnd = rowBase + chunkRows;
    if (n < rowEnd) rowEnd = n;
    if (rowEnd <= rowBase) return;
    const int tiles = (rowEnd - rowBase + kTileM - 1) / kTileM;
    const int L = net.nLevels;
    int* cursors = net.encCursor;  // [L] next unclaimed task per level; zero at launch, reset by k_query_mlp
    // (a static __shared__ variable would push the 227 KB dynamic allocation over the per-block limit)
    int& sNext = *reinterpret_cast<int*>(smem + smemBytes + (blockDim.x >> 5) * 2u * 32u * net.inputDims * 4u);
    int staged = -1;
    for (int level = blockIdx.x % L; level >= 0;) {
        LevelInfo lv;
        lv.scale = net.levelScale[level];
        lv.res = net.levelRes[level];
        lv.size = net.levelSize[level];
        lv.dense = net.levelDense[level] != 0;
        lv.pow2 = net.levelPow2[level] != 0;
        lv.mask = (lv.size - 1u) * (F * 2u);
        const unsigned char* gtable = reinterpret_cast<const unsigned char*>(net.grid + static_cast<size_t>(net.levelOffset[level]) * F);
        const uint32_t tableBytes = lv.size * F * 2u;
        const bool fits = tableBytes <= smemBytes && !(net.debug & 4);
        if (fits) {
            if (staged != level) {
                for (uint32_t i = threadIdx.x * 16u; i < tableBytes; i += blockDim.x * 16u)
                    *reinterpret_cast<uint4*>(smem + i) = __ldg(reinterpret_cast<const uint4*>(gtable + i));
                staged = level;
                __syncthreads();
            }
            encode_tasks<F, true>(net, lv, level, gtable, tc::s32(smem), in, rowBase, rowEnd, tiles, cursors + level, xt, ek,
                                  reinterpret_cast<float*>(smem + smemBytes));
        } else {
            encode_tasks<F, false>(net, lv, level, gtable, 0u, in, rowBase, rowEnd, tiles, cursors + level, xt, ek,
                                   reinterpret_cast<float*>(smem + smemBytes));
        }
        __syncthreads();  // every warp has found this level's cursor past the end
        // The levels do not cost the same (coarse levels gather a few shared cells, fine ones 24 random entries
        // per query): a CTA whose level is finished restages the table of the level furthest behind.
        if (threadIdx.x < 32) {  // one load per lane (L <= 16): the cursors arrive together, not one L2 round trip each
            const int at = static_cast<int>(threadIdx.x) < L ? *reinterpret_cast<volatile int*>(cursors + threadIdx.x) : tiles * 4;
            unsigned key = at < tiles * 4 ? static_cast<unsigned>(at) * 32u + threadIdx.x : 0xffffffffu;  // the cursors count tasks
            key = __reduce_min_sync(0xffffffffu, key);
            if (threadIdx.x == 0) sNext = key == 0xffffffffu ? -1 : static_cast<int>(key & 31u);
        }
        __syncthreads();
        level = sNext;
    }
}

// ---- MLP -----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t pack_leaky(uint32_t a, uint32_t b) { return leaky_pack2(__uint_as_float(a), __uint_as_float(b)); }
__device__ __forceinline__ uint32_t pack_plain(uint32_t a, uint32_t b) {
    const __half2 h = __floats2half2_rn(__uint_as_float(a), __uint_as_float(b));
    return *reinterpret_cast<const uint32_t*>(&h);
}

// Experiments (net.debug bit 4): CTA 0 stamps clock64 at protocol points into the rows after the
// last output row (the caller allocates 4096 extra rows): slot layout [who][event][4].
__device__ __forceinline__ void trace(bool on, __half* out, int n, int who, int ev, int field) {
    if (!on || ev >= 64) return;
    long long t;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t));
    reinterpret_cast<long long*>(out + static_cast<size_t>(n) * 16)[(who * 64 + ev) * 4 + field] = t;
}

// Barrier block at plan.offBar: [0] weights landed, [1] X full, [2] X empty, [3,4] MMA done (tile A, B),
// [5,6] epilogue done (tile A, B), then the TMEM base address.
template <int WIDTH>
__global__ void __launch_bounds__(kMlpThreads, 1)
    k_query_mlp(const __grid_constant__ NetDev net, int n, const int* dCount, int rowBase, int chunkRows,
                const unsigned char* __restrict__ xt, __half* __restrict__ out, const __grid_constant__ SplitPlan plan) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (dCount) {
        const int dn = *dCount;
        n = dn < n ? dn : n;
    }
    int rowEnd = rowBase + chunkRows;
    if (n < rowEnd) rowEnd = n;
    if (rowEnd <= rowBase) return;
    const int tiles = (rowEnd - rowBase + kTileM - 1) / kTileM;
    if (blockIdx.x == 0 && tid < net.nLevels) net.encCursor[tid] = 0;  // k_query_encode's tile cursors, for its next launch
    if (static_cast<int>(blockIdx.x) >= tiles) return;
    const bool tr = (net.debug & 16) && blockIdx.x == 0;
    const int nMine = (tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);

    const int EK = plan.ek, H = net.nHidden, LAYERS = H + 1;
#ifndef NRB_MLP_TS
#define NRB_MLP_TS 1
#endif
    // Hidden activations as the A operand in tensor memory (width 128): the epilogue writes the fp16 rows back with
    // tcgen05.st instead of st.shared + a generic-to-async proxy fence.  Columns: [0, 2W) the two accumulators,
    // [2W, 3W) the two activation tiles (W / 2 columns each, two K elements per column).
    constexpr bool kTS = NRB_MLP_TS && WIDTH == 128;
    constexpr uint32_t kTmemCols = kTS ? 512u : 2u * WIDTH;
    constexpr uint32_t kActBase = 2u * WIDTH, kActCols = WIDTH / 2u;
    // ... which frees the two hidden tiles' shared memory: the encoded tiles then arrive through a ring of two slots
    // (one otherwise), so a tile's load no longer waits for the layer-0 MMAs of the tile before it
#ifndef NRB_MLP_XSLOTS
#define NRB_MLP_XSLOTS 2
#endif
    constexpr int kXSlots = kTS ? NRB_MLP_XSLOTS : 1;
    const int nX = plan.xBytes <= plan.hBytes ? kXSlots : 1;  // the spare slots are the hidden tiles' (unused) space
    unsigned char* wsm = smem + plan.offW;
    unsigned char* X = smem + plan.offX;
    unsigned char* Hb = smem + plan.offH;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + plan.offBar);
    uint64_t* barW = bars + 0;
    // slot 0: bars 1 (full), 2 (empty); slots 1, 2: bars 8 / 9, 10 / 11
    auto barXFull = [&](int slot) { return slot == 0 ? bars + 1 : bars + 6 + 2 * slot; };
    auto barXEmpty = [&](int slot) { return slot == 0 ? bars + 2 : bars + 7 + 2 * slot; };
    auto xSlot = [&](int slot) { return slot == 0 ? X : Hb + static_cast<size_t>(slot - 1) * plan.hBytes; };
    uint64_t* barMma = bars + 3;   // [2]
    uint64_t* barEpi = bars + 5;   // [2]
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(bars + 7);

    if (tid == 0) {
        tc::mbar_init(barW, 1);
        for (int sl = 0; sl < nX; ++sl) {
            tc::mbar_init(barXFull(sl), 1);
            tc::mbar_init(barXEmpty(sl), 1);
        }
        tc::mbar_init(barMma + 0, 1);
        tc::mbar_init(barMma + 1, 1);
        tc::mbar_init(barEpi + 0, kEpiWarps);
        tc::mbar_init(barEpi + 1, kEpiWarps);
        tc::mbar_init_fence();
    }
    if (warp == 0) tc::tmem_alloc(tmemSlot, kTmemCols);
    tc::fence_before();
    __syncthreads();
    tc::fence_after();
    const uint32_t tmem = *tmemSlot;

    if (warp == 0) {
        // ===================== producer: weights once, then one encoded tile per X-slot turn ==========
        if (lane == 0) {
            tc::mbar_expect_tx(barW, plan.packedBytes);
            for (uint32_t o = 0; o < plan.packedBytes; o += 32768u) {
                const uint32_t b = plan.packedBytes - o < 32768u ? plan.packedBytes - o : 32768u;
                tc::bulk_g2s(wsm + o, static_cast<const unsigned char*>(net.packedW) + o, b, barW);
            }
            for (int k = 0; k < nMine; ++k) {
                const int tile = blockIdx.x + k * gridDim.x;
                const int sl = k % nX;
                tc::mbar_wait(barXEmpty(sl), ((k / nX) & 1) ^ 1);
                tc::mbar_expect_tx(barXFull(sl), plan.xBytes);
                tc::bulk_g2s(xSlot(sl), xt + static_cast<size_t>(tile) * plan.xBytes, plan.xBytes, barXFull(sl));
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer: two tiles in flight, offset by half a tile =================
        // The whole warp runs this loop convergently (uniform operands -> uniform registers); one
        // elected lane issues.  Tile context A runs layer ops at even turns, B at odd turns starting
        // half a tile later, so one context's epilogue runs under the other's MMAs.
        tc::mbar_wait(barW, 0);
        tc::fence_after();
        const uint32_t elected = tc::elect_one();
        const int nTA = (nMine + 1) >> 1, nTB = nMine >> 1;
        const int delay = LAYERS >> 1;
        const uint32_t wBase = tc::s32(wsm);
        const uint32_t idescW = tc::instr_desc(kTileM, WIDTH), idescL = tc::instr_desc(kTileM, 16);
        uint64_t dXs[kXSlots];
#pragma unroll
        for (int sl = 0; sl < kXSlots; ++sl) dXs[sl] = tc::smem_desc(tc::s32(xSlot(sl)), 2048u, 128u);
        const uint64_t dW0 = tc::smem_desc(wBase, 128u, static_cast<uint32_t>(EK) * 16u);
        const uint64_t dWh = tc::smem_desc(wBase + plan.w0Bytes, 128u, static_cast<uint32_t>(WIDTH) * 16u);
        const uint64_t dHA = tc::smem_desc(tc::s32(Hb), 2048u, 128u);
        const uint64_t dHB = tc::smem_desc(tc::s32(Hb + plan.hBytes), 2048u, 128u);
        const int ksteps0 = EK / 16;
        uint32_t xUse = 0;
        // one layer op of context c (compile-time), returns after the MMAs are queued and committed
        auto op = [&](auto cTag, int tk, int layer, uint32_t& epiPhase, int step) {
            constexpr int c = decltype(cTag)::value;
            const uint32_t acc = tmem + static_cast<uint32_t>(c * WIDTH);
            trace(tr && lane == 0, out, n, 0, step, 0);
            if (layer == 0) {
                if (tk > 0) {  // the previous tile's output epilogue has drained the accumulator
                    tc::mbar_wait(barEpi + c, epiPhase);
                    epiPhase ^= 1u;
                }
                const int sl = static_cast<int>(xUse % static_cast<uint32_t>(nX));
                tc::mbar_wait(barXFull(sl), (xUse / static_cast<uint32_t>(nX)) & 1u);
                ++xUse;
                tc::fence_after();
                trace(tr && lane == 0, out, n, 0, step, 1);
                uint64_t dX = dXs[0];
#pragma unroll
                for (int q = 1; q < kXSlots; ++q)
                    if (sl == q) dX = dXs[q];
                for (int k = 0; k < ksteps0; ++k)
                    tc::mma_f16_elect(elected, acc, dX + static_cast<uint64_t>(256u * k), dW0 + static_cast<uint64_t>(16u * k), idescW, k > 0);
                tc::commit_elect(elected, barXEmpty(sl));
                tc::commit_elect(elected, barMma + c);
            } else {
                tc::mbar_wait(barEpi + c, epiPhase);
                epiPhase ^= 1u;
                tc::fence_after();
                trace(tr && lane == 0, out, n, 0, step, 1);
                const uint64_t a0 = c == 0 ? dHA : dHB;
                const uint64_t b0 = dWh + static_cast<uint64_t>(static_cast<uint32_t>(layer - 1) * (WIDTH * WIDTH * 2u / 16u));
                if (kTS) {
                    const uint32_t aT = tmem + kActBase + static_cast<uint32_t>(c) * kActCols;
                    const uint32_t idesc = layer == H ? idescL : idescW;
#pragma unroll
                    for (int k = 0; k < WIDTH / 16; ++k)
                        tc::mma_f16_ts_elect(elected, acc, aT + 8u * k, b0 + static_cast<uint64_t>(16u * k), idesc, k > 0);
                } else if (layer == H) {
#pragma unroll
                    for (int k = 0; k < WIDTH / 16; ++k)
                        tc::mma_f16_elect(elected, acc, a0 + static_cast<uint64_t>(256u * k), b0 + static_cast<uint64_t>(16u * k), idescL, k > 0);
                } else {
#pragma unroll
                    for (int k = 0; k < WIDTH / 16; ++k)
                        tc::mma_f16_elect(elected, acc, a0 + static_cast<uint64_t>(256u * k), b0 + static_cast<uint64_t>(16u * k), idescW, k > 0);
                }
                tc::commit_elect(elected, barMma + c);
            }
            trace(tr && lane == 0, out, n, 0, step, 2);
        };
        uint32_t phA = 0, phB = 0;
        int tkA = 0, layA = 0, tkB = 0, layB = 0;
        for (int turn = 0; tkA < nTA || tkB < nTB; ++turn) {
            if (tkA < nTA) {
                op(std::integral_constant<int, 0>{}, tkA, layA, phA, 2 * turn);
                if (++layA == LAYERS) { layA = 0; ++tkA; }
            }
            if (tkB < nTB && turn >= delay) {
                op(std::integral_constant<int, 1>{}, tkB, layB, phB, 2 * turn + 1);
                if (++layB == LAYERS) { layB = 0; ++tkB; }
            }
        }
    } else {
        // ===================== epilogue warps: TMEM -> LeakyReLU -> fp16 -> next layer's A operand ====
        const int c = (warp - 2) / kEpiWarps;       // tile context
        const int half = ((warp - 2) % kEpiWarps) >> 2;  // which half of the columns
        const int quarter = warp & 3;               // the TMEM lane quarter this warp may read
        const int row = quarter * 32 + lane;
        const uint32_t acc = tmem + (static_cast<uint32_t>(quarter * 32) << 16) + static_cast<uint32_t>(c * WIDTH);
        unsigned char* Hc = Hb + c * plan.hBytes;
        const int nT = c == 0 ? (nMine + 1) >> 1 : nMine >> 1;
        constexpr int kCols = WIDTH / 2;            // columns per thread per hidden layer
        uint32_t phase = 0;
        for (int tk = 0; tk < nT; ++tk) {
            const int tile = blockIdx.x + (2 * tk + c) * gridDim.x;
            const int q = rowBase + tile * kTileM + row;
            for (int layer = 0; layer < H; ++layer) {
                const bool trE = tr && lane == 0 && ((warp - 2) % kEpiWarps) == 0;
                const int ev = tk * (H + 1) + layer;
                trace(trE, out, n, 1 + c, ev, 0);
                tc::mbar_wait(barMma + c, phase);
                phase ^= 1u;
                tc::fence_after();
                trace(trE, out, n, 1 + c, ev, 1);
                uint32_t v[kCols / 32][32];
#pragma unroll
                for (int b = 0; b < kCols / 32; ++b) tc::ld32(acc + half * kCols + b * 32, v[b]);
                tc::wait_ld();
                trace(trE, out, n, 1 + c, ev, 2);
                if constexpr (kTS) {
                    uint32_t pk[32];
#pragma unroll
                    for (int b = 0; b < kCols / 32; ++b)
#pragma unroll
                        for (int j = 0; j < 16; ++j) pk[b * 16 + j] = pack_leaky(v[b][2 * j], v[b][2 * j + 1]);
                    tc::st32(tmem + (static_cast<uint32_t>(quarter * 32) << 16) + kActBase + static_cast<uint32_t>(c) * kActCols +
                                 static_cast<uint32_t>(half) * 32u, pk);
                    tc::wait_st();
                } else {
#pragma unroll
                for (int b = 0; b < kCols / 32; ++b)
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        uint4 o;
                        o.x = pack_leaky(v[b][8 * j + 0], v[b][8 * j + 1]);
                        o.y = pack_leaky(v[b][8 * j + 2], v[b][8 * j + 3]);
                        o.z = pack_leaky(v[b][8 * j + 4], v[b][8 * j + 5]);
                        o.w = pack_leaky(v[b][8 * j + 6], v[b][8 * j + 7]);
                        *reinterpret_cast<uint4*>(Hc + ((((half * kCols) >> 3) + b * 4 + j) * kTileM + row) * 16) = o;
                    }
                tc::proxy_fence();
                }
                tc::fence_before();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(barEpi + c);
                trace(trE, out, n, 1 + c, ev, 3);
            }
            tc::mbar_wait(barMma + c, phase);
            phase ^= 1u;
            tc::fence_after();
            {
                uint32_t v[8];
                tc::ld8(acc + half * 8, v);
                tc::wait_ld();
                if (q < rowEnd) {
                    uint4 o;
                    o.x = pack_plain(v[0], v[1]); o.y = pack_plain(v[2], v[3]);
                    o.z = pack_plain(v[4], v[5]); o.w = pack_plain(v[6], v[7]);
                    reinterpret_cast<uint4*>(out + static_cast<size_t>(q) * 16)[half] = o;
                }
            }
            tc::fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(barEpi + c);
        }
    }
    tc::fence_before();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, kTmemCols);
}

}  // namespace

// ---- host side -----------------------------------------------------------------------------------
bool split_query_supported(const NetDev& net) {
    if (net.width != 64 && net.width != 128) return false;
    if (net.nFeatures != 2 && net.nFeatures != 4) return false;
    if (net.nHidden < 1 || net.shDegree > 4 || net.nLevels < 1 || net.nLevels > kMaxLevels) return false;
    if (((net.nPoints * net.nLevels * net.nFeatures) & 7) != 0) return false;
    const SplitPlan p = make_plan(net);
    return p.mlpSmem <= kSmemMax && p.encSmem + p.encRing <= kSmemMax;
}

size_t split_packed_bytes(const NetDev& net) { return make_plan(net).packedBytes; }

size_t split_xt_bytes(const NetDev& net, size_t rows) {
    const size_t tiles = (rows + kTileM - 1) / kTileM;
    return tiles * make_plan(net).xBytes;
}

int split_prepack(const NetDev& net, void* packedW, cudaStream_t s) {
    const SplitPlan p = make_plan(net);
    k_pack_weights<<<64, 256, 0, s>>>(net, static_cast<unsigned char*>(packedW), p.ek);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

void init_query_split_kernels() {
    cudaFuncSetAttribute(k_query_encode<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemMax));
    cudaFuncSetAttribute(k_query_encode<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemMax));
    cudaFuncSetAttribute(k_query_mlp<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemMax));
    cudaFuncSetAttribute(k_query_mlp<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemMax));
}

int launch_split_query(const NetDev& net, const float* in, size_t n, const int* dCount, __half* out, cudaStream_t s) {
    if (n == 0) return 0;
    if (!net.packedW || !net.xt || net.xtRows < kTileM || !net.encCursor) return -1;
    const SplitPlan p = make_plan(net);
    const int sms = sm_count();
    int chunkRows = static_cast<int>(net.xtRows / kTileM * kTileM);
    if ((net.debug & 8) && chunkRows > 1024) chunkRows = 1024;  // tests: exercise the chunk walk on small batches
    {
        const int lg = (net.debug >> 8) & 31;  // experiments: NRB_OPT_QUERY_KERNEL = 3 | (log2 rows << 16) walks the batch in smaller chunks
        if (lg >= 10 && chunkRows > (1 << lg)) chunkRows = 1 << lg;
    }
    int launches = 0;
    for (size_t base = 0; base < n; base += chunkRows) {
        const size_t rows = n - base < static_cast<size_t>(chunkRows) ? n - base : chunkRows;
        const int tiles = static_cast<int>((rows + kTileM - 1) / kTileM);
        // one CTA per SM, never more than four tiles' worth of CTAs per level
        const int want = net.nLevels * ((tiles + 3) / 4);
        const int egrid = want < sms ? want : sms;
        unsigned char* xt = static_cast<unsigned char*>(net.xt);
        if (net.nFeatures == 4)
            k_query_encode<4><<<egrid, kEncThreads, p.encSmem + p.encRing + 16u, s>>>(net, in, static_cast<int>(n), dCount, static_cast<int>(base),
                                                                    chunkRows, xt, p.ek, p.encSmem);
        else
            k_query_encode<2><<<egrid, kEncThreads, p.encSmem + p.encRing + 16u, s>>>(net, in, static_cast<int>(n), dCount, static_cast<int>(base),
                                                                    chunkRows, xt, p.ek, p.encSmem);
        NRB_LAUNCHED("k_query_encode", s);
        const int mgrid = tiles < sms ? tiles : sms;
        if (net.width == 128)
            k_query_mlp<128><<<mgrid, kMlpThreads, p.mlpSmem, s>>>(net, static_cast<int>(n), dCount, static_cast<int>(base), chunkRows,
                                                                  xt, out, p);
        else
            k_query_mlp<64><<<mgrid, kMlpThreads, p.mlpSmem, s>>>(net, static_cast<int>(n), dCount, static_cast<int>(base), chunkRows,
                                                                 xt, out, p);
        NRB_LAUNCHED("k_query_mlp", s);
        launches += 2;
    }
    return launches;
}

// ---- the split query on static scratch (nrb_neural_query_static: no allocation inside the call) ----------
// The encoded tiles and the level cursors live in ONE file-scope device array instead of the renderer's cudaMalloc'ed
// buffers: 16 384 rows of at most 512 bytes (EK <= 256 encoded fp16 values), then the cursors.  Every GPU that loads
// the library holds these 8 MiB + 256 B until exit.  All calls on a device share the array: calls that overlap on the
// GPU (two streams, two renderers) race on it.
constexpr size_t kStaticRows = 16384;
constexpr size_t kStaticRowBytes = 512;
constexpr size_t kStaticXtBytes = kStaticRows * kStaticRowBytes;
static __device__ __align__(256) unsigned char g_staticScratch[kStaticXtBytes + 256];

size_t split_static_max_rows() { return kStaticRows; }

// g_staticScratch is a host-side stand-in here, not a device address: ask the runtime once per device, with that
// device current (the caller's cudaSetDevice), and keep the answer.
static cudaError_t static_scratch_address(unsigned char** out) {
    static unsigned char* cached[64] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!cached[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_staticScratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        cached[dev] = static_cast<unsigned char*>(p);
    }
    *out = cached[dev];
    return cudaSuccess;
}

int launch_split_query_static(const NetDev& net, const float* in, size_t n, __half* out, cudaStream_t s, const char** why) {
    *why = nullptr;
    if (n == 0) return 0;
    if (!net.packedW || !split_query_supported(net)) {
        *why = "the static-scratch query needs a network shape the split query supports (width 64/128, 2 or 4 features)";
        return -1;
    }
    if (n > kStaticRows || split_xt_bytes(net, n) > kStaticXtBytes) {
        *why = "too many queries for the static scratch (at most 16384 per call, encoded width <= 256)";
        return -1;
    }
    unsigned char* base = nullptr;
    const cudaError_t e = static_scratch_address(&base);
    if (e != cudaSuccess) {
        *why = cudaGetErrorString(e);
        return -2;
    }
    NetDev tmp = net;
    tmp.debug = 0;
    tmp.xt = base;
    tmp.xtRows = static_cast<uint32_t>((n + kTileM - 1) / kTileM * kTileM);
    tmp.encCursor = reinterpret_cast<int*>(base + kStaticXtBytes);
    // the cursors must be zero at launch; k_query_mlp leaves them so, but a call that failed half-way would not have
    if (cudaMemsetAsync(tmp.encCursor, 0, kMaxLevels * sizeof(int), s) != cudaSuccess) {
        *why = "cudaMemsetAsync on the static scratch failed";
        return -2;
    }
    return launch_split_query(tmp, in, n, nullptr, out, s);
}

}  // namespace nrb
