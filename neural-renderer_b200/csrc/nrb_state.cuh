// nrb_state.cuh — per-element wavefront state of one renderer, 16-byte records in SoA arrays.
//
// The reference keeps 384 B per element in ~45 scalar/triplet AoS arrays sized W*H*spp
// (src/cuda_renderer_neural.cu:2890-2952) and runs every kernel over all of them.  Here state is
// packed into float4 records (one LDG/STG.128 each), stages communicate through compacted index
// queues whose counts stay on the device, and stages that the reference separates only because of
// its kernel-per-step structure are fused so most of this state never leaves registers.
// Element index e = localPixel + s * localPixels (for world == 1 this is the reference's
// sampleIdx = pixelIdx + s * pixelCount, which makes the capture buffers directly comparable).
#pragma once

#include <cuda_fp16.h>

#include "nrb_scene.cuh"

namespace nrb {

// A surface record: what the reference spreads over hitPositions_/hitNormals_/hitColors_/
// hitMaterialParams_/hitFlags_/hitDistances_.
struct SurfRec {
    float4* P;  // position.xyz, flag (int bits)
    float4* N;  // normal.xyz, neural distance
    float4* C;  // albedo.xyz, metallic
    float2* M;  // roughness, specular
};

struct State {
    SurfRec rec[3];     // 0 = primary hit, 1 = bounce hit, 2 = additional-mesh hit
    float4* T;          // throughput.xyz, -
    float4* L;          // radiance.xyz, -
    float4* rayO;       // bounce origin.xyz, pdf
    float4* rayD;       // bounce direction.xyz, distance carried by the sampled vertex
    float4* rayW;       // BRDF weight.xyz, -
    float4* prevD;      // incoming direction of the current vertex (bounce >= 2)
    float4* segE;       // shell entry position.xyz, accumT
    float4* segD;       // segment direction.xyz, -
    float4* segX;       // segment exit position.xyz, outerExitT
    float2* segI;       // innerEnterT, innerHit flag (int bits)
    int* queue[4];      // 0/1: path queues (ping-pong), 2/3: segment queues (ping-pong)
    int* counts;        // [0..3] queue counts, [4..] scratch; cleared by the host per stage
    const int* tileOrder;   // primary-pass 8x4 tiles, most expensive first (from tileCost of an earlier frame), or null
    unsigned int* tileCost; // clocks each tile's warp spent in the primary pass of this frame
    unsigned long long* stats;  // [0] ray segments traced, [1] neural queries, [2] path vertices,
                                // [3] node pairs fetched, [4] triangles fetched
    float* netIn;       // [slot][inputDims] fp32, slot = position in the segment queue
    __half* netOut;     // [slot][16] fp16
    float* accum;       // 3 per local pixel (the reference's accum_)
    uchar4* pixels;     // full frame, uchar4 RGBA
};

__device__ __forceinline__ void rec_store(const SurfRec& r, int e, V3 p, int flag, V3 n, float dist, V3 c, V3 mp) {
    r.P[e] = make_float4(p.x, p.y, p.z, __int_as_float(flag));
    r.N[e] = make_float4(n.x, n.y, n.z, dist);
    r.C[e] = make_float4(c.x, c.y, c.z, mp.x);
    r.M[e] = make_float2(mp.y, mp.z);
}

// Warp-aggregated queue append (one atomic per warp instead of the reference's one per element,
// src/cuda_renderer_neural.cu:1529-1541).  Must be called by all active lanes with a predicate.
__device__ __forceinline__ void queue_push(int* queue, int* count, bool pred, int value) {
    const unsigned mask = __activemask();
    const unsigned votes = __ballot_sync(mask, pred);
    if (votes == 0) return;
    const int lane = threadIdx.x & 31;
    const int leader = __ffs(votes) - 1;
    int base = 0;
    if (lane == leader) base = atomicAdd(count, __popc(votes));
    base = __shfl_sync(mask, base, leader);
    if (pred) queue[base + __popc(votes & ((1u << lane) - 1))] = value;
}

}  // namespace nrb
