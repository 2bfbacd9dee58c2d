// nrb_frame_dev.cuh — device helpers shared by the frame kernels (ground-truth and neural paths).
#pragma once

#include "nrb_frame.cuh"
#include "nrb_shade.cuh"
#include "nrb_trace.cuh"

namespace nrb {

constexpr int kTileThreads = 128;  // 4 warps, each an 8x4 pixel tile -> a block covers 32x4 pixels
constexpr int kQueueThreads = 128;

struct ElemId {
    int e, x, gy, s, pixel;  // element, column, GLOBAL row, sample, global pixel index
    bool valid;
};

// Primary-ray kernels: warp = 8x4 pixel tile (coherent rays share BVH nodes in L1).
__device__ __forceinline__ ElemId tile_elem(const FrameParams& fp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    ElemId id;
    id.x = blockIdx.x * 32 + warp * 8 + (lane & 7);
    const int ly = blockIdx.y * 4 + (lane >> 3);
    id.s = blockIdx.z;
    id.valid = id.x < fp.width && ly < fp.localHeight;
    id.gy = global_row(fp, ly);
    id.pixel = id.gy * fp.width + id.x;
    id.e = ly * fp.width + id.x + id.s * fp.localPixels;
    return id;
}

// The same 8x4 tiles handed out one per warp from a device cursor (persistent grids: the resident warps of every SM
// stay busy until the frame's last tile, and tiles of sky pixels cost no CTA launch).  tile < 0: none left.
__device__ __forceinline__ int tile_count(const FrameParams& fp) {
    return ((fp.width + 7) >> 3) * ((fp.localHeight + 3) >> 2) * fp.spp;
}
__device__ __forceinline__ int next_tile(int* cursor, int total) {
    int t = 0;
    if ((threadIdx.x & 31) == 0) t = atomicAdd(cursor, 1);
    t = __shfl_sync(0xffffffffu, t, 0);
    return t < total ? t : -1;
}
__device__ __forceinline__ ElemId tile_elem_at(const FrameParams& fp, int tile) {
    const int lane = threadIdx.x & 31;
    const int tx = (fp.width + 7) >> 3, ty = (fp.localHeight + 3) >> 2;
    ElemId id;
    id.s = tile / (tx * ty);
    const int r = tile - id.s * (tx * ty);
    const int by = r / tx, bx = r - by * tx;
    id.x = bx * 8 + (lane & 7);
    const int ly = by * 4 + (lane >> 3);
    id.valid = id.x < fp.width && ly < fp.localHeight;
    id.gy = global_row(fp, ly);
    id.pixel = id.gy * fp.width + id.x;
    id.e = ly * fp.width + id.x + id.s * fp.localPixels;
    return id;
}

__device__ __forceinline__ ElemId queue_elem(const FrameParams& fp, int e) {
    ElemId id;
    id.e = e;
    id.s = e / fp.localPixels;
    const int lp = e - id.s * fp.localPixels;
    const int ly = lp / fp.width;
    id.x = lp - ly * fp.width;
    id.gy = global_row(fp, ly);
    id.pixel = id.gy * fp.width + id.x;
    id.valid = true;
    return id;
}

// First queue position of this thread in a grid-sized-stride loop.  Long queues: consecutive positions go to
// consecutive lanes (neighbouring pixels share a warp).  A queue with fewer than four elements per warp - the
// later bounces, the later segment iterations: a few thousand survivors - is spread one element per warp and
// lane row (lane L of warp w starts at L * warps + w): such launches are latency-bound, and a warp that traces
// r rays in lock step takes, phase by phase, as long as its slowest lane, so r should be as small as possible.
// `perWarp`: spread when the queue has at most this many elements per warp; stages that trace (divergent, no
// coalescing to lose) pass 32 - every queue that fits the grid once is spread over all its warps.
__device__ __forceinline__ int queue_start(int n, int perWarp = 4) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int warps = (gridDim.x * blockDim.x) >> 5;
    if (n > perWarp * warps) return tid;
    return (threadIdx.x & 31) * warps + (tid >> 5);
}

__device__ __forceinline__ void stat_flush(unsigned long long* slot, unsigned int v) {
    // all threads of the block reach this point
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(slot, static_cast<unsigned long long>(v));
}

__device__ __forceinline__ V3 f4v(float4 a) { return v3(a.x, a.y, a.z); }
__device__ __forceinline__ V3 global_mparams(const MaterialDev& g) {
    return v3(g.metallic.value, g.roughness.value, g.specular.value);
}

// Trace one mesh and build the reference's hit record (the tail shared by :244-296, :556-618,
// :1128-1188, :1235-1302).
template <int MODE, bool INLINE = false>
__device__ __forceinline__ bool hit_record(const Frame& f, int slot, V3 o, V3 d, V3& pos, V3& nrm, V3& alb, V3& mp) {
    const MeshDev& m = f.mesh[slot];
    // INLINE: kernels whose only traversal this is (ground-truth primary / bounce) inline the loop - a called copy
    // moves the global-memory descriptor to a uniform register before every load
    const Hit h = INLINE ? trace_dyn_inline(target_of(m), MODE, o, d) : trace_dyn(target_of(m), MODE, o, d);
    if (h.t >= 1e30f) {
        mp = global_mparams(f.gmat);
        return false;
    }
    const Surface s = surface_at(m, h);
    pos = s.position;
    nrm = s.normal;
    resolve_surface_material(m, s, f.gmat, alb, mp);
    return true;
}

// hit_record for a converged warp (every lane calls, `valid` lanes hold a ray): the traversal is trace_sync, the same walk
// reconverged every few steps.  Same hit, same record.
template <int MODE>
__device__ __forceinline__ bool hit_record_sync(const Frame& f, int slot, bool valid, V3 o, V3 d, V3& pos, V3& nrm, V3& alb,
                                                V3& mp) {
    const MeshDev& m = f.mesh[slot];
    const Hit h = trace_sync(target_of(m), MODE, valid, o, d);
    if (!valid || h.t >= 1e30f) {
        mp = global_mparams(f.gmat);
        return false;
    }
    const Surface s = surface_at(m, h);
    pos = s.position;
    nrm = s.normal;
    resolve_surface_material(m, s, f.gmat, alb, mp);
    return true;
}

// One bounce vertex: sampleBounceDirectionsKernel for a live path.  Returns false when the
// reference deactivates the path (pdf <= 0, NaN/Inf weight).
__device__ __forceinline__ bool sample_vertex(const Frame& f, const ElemId& id, V3 hitPos, V3 normal, V3 albedo,
                                              V3 mp, bool firstBounce, V3 prevDir, V3& org, V3& wi, V3& weight,
                                              float& pdf) {
    const FrameParams& fp = f.fp;
    uint32_t rng = rng_seed(id.pixel, fp.sampleOffset, id.s);
    const float nl = length(normal);
    normal = nl > 0.0f ? normal / nl : v3(0.0f, 1.0f, 0.0f);
    const V3 incoming = firstBounce ? primary_direction(fp, id.x, id.gy, rng) : prevDir;
    if (dot(normal, incoming) > 0.0f) normal = normal * -1.0f;
    const V3 wo = incoming * -1.0f;
    V3 tg, bt;
    tangent_frame(normal, tg, bt);
    const Bsdf b = bsdf_from(f.gmat, albedo, mp);
    const float u1 = rng_next(rng), u2 = rng_next(rng), u3 = rng_next(rng);
    wi = disney_sample3(b, normal, wo, tg, bt, u1, u2, u3, pdf);
    if (pdf <= 0.0f) return false;
    const V3 fr = disney_f(b, normal, wo, wi, tg, bt);
    const float ct = fabsf(dot(normal, wi));
    weight = fr * (ct / pdf);
    if (isnan(weight.x) || isnan(weight.y) || isnan(weight.z) || isinf(weight.x) || isinf(weight.y) ||
        isinf(weight.z))
        return false;
    org = hitPos + normal * (fp.sceneScale * 1e-6f);
    return true;
}

// integrateBounceKernel for a live path whose bounce ray has been resolved (hit or miss).
// Returns true when the path continues; T and L are updated in place (L only on a miss).
__device__ __forceinline__ bool integrate_vertex(const Frame& f, const ElemId& id, int bounce, bool hit, V3 wi,
                                                 V3 weight, V3& T, V3& L, bool& touchedL) {
    const FrameParams& fp = f.fp;
    T = hmul(T, weight);
    touchedL = false;
    if (!hit) {
        L = L + hmul(T, env_clamped(f.env, wi, fp.maxRadiance));
        touchedL = true;
        return false;
    }
    if (bounce >= fp.maxBounces) return false;
    const float p = fminf(1.0f, fmaxf(T.x, fmaxf(T.y, T.z)));
    uint32_t rng = rng_seed(id.pixel, fp.sampleOffset + static_cast<uint32_t>(bounce), id.s);
    if (rng_next(rng) >= p) return false;
    T = T * (1.0f / p);
    return true;
}

__device__ __forceinline__ uchar4 to_srgb8(V3 c) {
    return make_uchar4(static_cast<unsigned char>(srgb_encode(c.x) * 255.0f),
                       static_cast<unsigned char>(srgb_encode(c.y) * 255.0f),
                       static_cast<unsigned char>(srgb_encode(c.z) * 255.0f), 255);
}

inline int queue_grid(int maxElems) {
    // persistent-style grid: enough CTAs to fill every SM with 8 resident CTAs, no more than needed
    const int need = (maxElems + kQueueThreads - 1) / kQueueThreads;
    const int cap = sm_count() * 8;
    return need < 1 ? 1 : (need < cap ? need : cap);
}

}  // namespace nrb
