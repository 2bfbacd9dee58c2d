// nrb_segments.cu — frame kernels of the neural path: shell-segment state machine, neural hit
// decoding, hybrid additional mesh, and the bounce stages around them.  -fmad=false.
//
// Reference (src/cuda_renderer_neural.cu) -> here:
//   traceOuterShellEntryKernel :720-767                                   -> k_entry_camera
//   sampleBounceDirectionsKernel :357-499 + checkBounceEarlyTermination-
//     Kernel :503-553 + traceOuterShellEntryFromRaysKernel :773-852       -> k_bounce_begin
//   traceSegmentExitsKernel :858-930 + buildSegmentNeuralInputsKernel
//     :937-1001                                                           -> k_seg_exits
//   applySegmentNeuralOutputKernel :1007-1123 + prepareNextIteration-
//     Kernel :1456-1524 + compactInputsKernel :1529-1541                  -> k_seg_apply
//   traceAdditionalMeshPrimaryRaysKernel :1128-1188 + selectClosestPrimary-
//     HitKernel :1307-1378 + initializePathStateKernel :303-354           -> k_primary_finish
//   traceAdditionalMeshRaysKernel :1235-1302 + selectClosestHitKernel
//     :1383-1450 + integrateBounceKernel :621-675                         -> k_bounce_finish
// The host loop of traceNeuralSegmentsForRays (:1810-2083) reads the active count back to the
// host after every compaction; here the counts never leave the device: every stage is launched
// with a fixed persistent grid and loops over the count it finds in device memory.
#include <cstdio>
#include <cstring>

#include "nrb_frame_dev.cuh"
#include "nrb_neural.cuh"
#include "nrb_query_warp.cuh"

namespace nrb {

constexpr float kSegEps = 1e-8f;       // kSegmentEpsilon, :715
constexpr int kMaxSegIters = 10;       // kMaxSegmentIterations, :714

__device__ __forceinline__ V3 shell_norm(const FrameParams& fp, V3 p) {
    return v3((p.x - fp.outerMin.x) * fp.outerInvExtent.x, (p.y - fp.outerMin.y) * fp.outerInvExtent.y,
              (p.z - fp.outerMin.z) * fp.outerInvExtent.z);
}

__device__ __forceinline__ void rec_clear(const SurfRec& r, int e) {
    r.P[e] = make_float4(0.f, 0.f, 0.f, 0.f);
    r.N[e] = make_float4(0.f, 0.f, 0.f, 0.f);
    r.C[e] = make_float4(0.f, 0.f, 0.f, 0.f);
    r.M[e] = make_float2(0.f, 0.f);
}

__global__ void __launch_bounds__(kTileThreads) k_entry_camera(const __grid_constant__ Frame f) {
    const FrameParams& fp = f.fp;
    const ElemId id = tile_elem(fp);
    bool active = false;
    if (id.valid) {
        uint32_t rng = rng_seed(id.pixel, fp.sampleOffset, id.s);
        const V3 dir = primary_direction(fp, id.x, id.gy, rng);
        const Hit h = trace_closest<TRACE_FORWARD>(f.mesh[NRB_SLOT_OUTER], fp.camPos, dir);
        rec_clear(f.st.rec[0], id.e);
        f.st.segD[id.e] = make_float4(dir.x, dir.y, dir.z, 0.0f);
        if (h.t < 1e30f) {
            const V3 p = fp.camPos + dir * h.t;
            f.st.segE[id.e] = make_float4(p.x, p.y, p.z, h.t);
            active = true;
        }
    }
    queue_push(f.st.queue[2], f.st.counts + 2, active, id.e);
    stat_flush(f.st.stats + 0, id.valid ? 1u : 0u);
}

// One iteration, first half: exits + network inputs for every ray in the segment queue.
// traceSegmentExitsKernel + buildSegmentNeuralInputsKernel for one ray: exits into segX / segI, the D network
// inputs into dst.
// `pass`: -1 = both traces in this lane; 0 / 1 = this lane traces only the outer / the inner shell and its
// partner lane (lane ^ 1, same ray) the other one - sparse launches, where the latency of one lane's chain of
// traces is all there is.  All 32 lanes call (valid = false: no ray).
__device__ __forceinline__ void seg_exit_one(const Frame& f, int e, float* dst, int pass = -1, bool valid = true) {
    const FrameParams& fp = f.fp;
    const int D = fp.inputDims;
    V3 entry = v3(0.f, 0.f, 0.f), dir = v3(0.f, 0.f, 1.f);
    if (valid) {
        entry = f4v(f.st.segE[e]);
        dir = f4v(f.st.segD[e]);
    }
    const V3 shifted = entry + dir * kSegEps;
    // inner shell ANY, then outer shell BACKWARD: passes through ONE inlined copy of the traversal loop (k = 1 inner,
    // k = 0 outer).  With both in this lane the outer trace only looks as far as the inner hit: the exit distance is
    // used as a number only when the ray leaves the shell before it reaches the inner shell, otherwise only as
    // "beyond the inner hit" (stored as 1e30).  What this cannot tell apart is a ray with an inner hit and NO back
    // face of the outer shell at all - a leak through an edge of a closed shell - for which the reference falls back
    // to exitT = 1e-8 (:899-905); NRB_OPT_LITERAL_BOUNCE_TRACES traces to the end, as the reference does.
    float tOuter = 1e30f, tInner = 1e30f;
    bool beyond = false;
    const int k0 = !valid ? 2 : (pass < 0 ? 1 : pass), k1 = pass < 0 ? 0 : pass;
#pragma unroll 1
    for (int k = k0; k >= k1 && k < 2; --k) {
        const MeshDev& m = f.mesh[k == 0 ? NRB_SLOT_OUTER : NRB_SLOT_INNER];
        float tMax = 1e30f;
        if (k == 0 && pass < 0 && tInner < 1e30f && !f.literalSegExits) tMax = __uint_as_float(__float_as_uint(tInner) + 1u);
        const Hit h = trace_dyn_inline(target_of(m), k == 0 ? TRACE_BACKWARD : TRACE_ANY, shifted, dir, tMax);
        if (k == 0) {
            beyond = tMax < 1e30f && !(h.t < tMax);
            tOuter = beyond ? 1e30f : h.t;
        } else {
            tInner = h.t;
        }
    }
    if (pass >= 0) {
        const float mine = pass == 0 ? tOuter : tInner;
        const float other = __shfl_xor_sync(0xffffffffu, mine, 1);
        if (pass == 0) tInner = other; else tOuter = other;
        if (!valid || pass != 0) return;  // the even lane of a couple writes
    }
    const float exitT = tOuter < 1e30f ? tOuter : (beyond ? 1e30f : kSegEps);
    const bool hitInner = tInner < 1e30f;
    const float innerT = hitInner ? tInner : 1e30f;
    const bool innerFirst = hitInner && (innerT < exitT);
    const V3 exitPos = innerFirst ? shifted + dir * innerT : shifted + dir * exitT;
    f.st.segX[e] = make_float4(exitPos.x, exitPos.y, exitPos.z, exitT);
    f.st.segI[e] = make_float2(innerT, __int_as_float(hitInner ? 1 : 0));
    const V3 a = shell_norm(fp, shifted), b = shell_norm(fp, exitPos);
    dst[0] = a.x; dst[1] = a.y; dst[2] = a.z;
    dst[3] = b.x; dst[4] = b.y; dst[5] = b.z;
    int o = 6;
    if (fp.nPoints == 3) {
        const V3 c = shell_norm(fp, (shifted + exitPos) * 0.5f);
        dst[6] = c.x; dst[7] = c.y; dst[8] = c.z;
        o = 9;
    }
    if (D >= o + 3) {
        dst[o] = (dir.x + 1.0f) * 0.5f;
        dst[o + 1] = (dir.y + 1.0f) * 0.5f;
        dst[o + 2] = (dir.z + 1.0f) * 0.5f;
    }
}

// `dynamic`: warps take 32 consecutive queue positions at a time from the cursor at counts[9] (reset by the
// host with the bounce cursor) instead of a fixed stride: the per-warp cost of 32 rays varies, and with the
// static partition the launch ended on its slowest warps at 36 % resident warps (profiles/r1_final).
__device__ __forceinline__ void seg_exits_body(const Frame& f, int qin, bool dynamic = false) {
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[qin];
    const int D = fp.inputDims;
    const int total = gridDim.x * blockDim.x;
    // short queues: one ray per warp (queue_start); up to half a ray per lane: the (ray, trace) couples below
    dynamic = dynamic && n > (total >> 5) * 4 && 2 * n > total;
    if (!dynamic && 2 * n <= total) {
        // sparse: one (ray, trace) task per lane, the couples of a queue spread over all warps
        const int warps = total >> 5, lane = threadIdx.x & 31;
        const int i = (lane >> 1) * warps + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
        const bool valid = i < n;
        const int e = valid ? f.st.queue[qin][i] : 0;
        seg_exit_one(f, e, f.st.netIn + static_cast<size_t>(valid ? i : 0) * D, lane & 1, valid);
    } else {
    int* cursor = f.st.counts + 9;
    int chunk = 0;
    for (int i = dynamic ? n : queue_start(n, 32);; i += total) {
        if (dynamic) {
            if ((threadIdx.x & 31) == 0) chunk = atomicAdd(cursor, 32);
            chunk = __shfl_sync(0xffffffffu, chunk, 0);
            if (chunk >= n) break;
            i = chunk + (threadIdx.x & 31);
            if (i >= n) continue;
        } else if (i >= n) {
            break;
        }
        seg_exit_one(f, f.st.queue[qin][i], f.st.netIn + static_cast<size_t>(i) * D);
    }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(f.st.stats + 0, 2ull * static_cast<unsigned long long>(n));
        atomicAdd(f.st.stats + 1, static_cast<unsigned long long>(n));
    }
}
__global__ void __launch_bounds__(kQueueThreads) k_seg_exits(const __grid_constant__ Frame f, int qin, int dynamic) {
    seg_exits_body(f, qin, dynamic != 0);
}

// applySegmentNeuralOutputKernel + prepareNextIterationKernel for one ray, in two halves.  seg_apply_decode:
// decode the 16 fp16 outputs and record the hit; returns -1 when the ray hit, else the inner-hit flag (0 / 1) of a
// ray that left the segment and must look for its re-entry.  seg_apply_reenter: that trace; true when the ray goes
// on to another segment.
__device__ __forceinline__ int seg_apply_decode(const Frame& f, int e, uint4 raw, const SurfRec& rec) {
    const FrameParams& fp = f.fp;
    const float2 o01 = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
    const float2 o23 = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
    const float2 o45 = __half22float2(*reinterpret_cast<const __half2*>(&raw.z));
    const float2 o67 = __half22float2(*reinterpret_cast<const __half2*>(&raw.w));
    const float2 I = f.st.segI[e];
    const float innerT = I.x, exitT = f.st.segX[e].w;
    const bool innerHit = __float_as_int(I.y) != 0;
    const bool neuralHit = o01.x >= 0.0f;
    if (neuralHit || (innerHit && innerT < exitT)) {
        const V3 entry = f4v(f.st.segE[e]), dir = f4v(f.st.segD[e]);
        const V3 shifted = entry + dir * kSegEps;
        const float distance = o01.y;
        float d = distance * 1.0f;
        if (innerT > 0.0f) d = fminf(d, innerT - kSegEps);
        const V3 pos = shifted + dir * d;
        V3 nrm = v3(o23.x, o23.y, o45.x);
        const float nl = length(nrm);
        nrm = nl > 1e-6f ? nrm / nl : v3(0.0f, 1.0f, 0.0f);
        const V3 col = fp.useConstantNeuralColor ? fp.constantNeuralColor
                                                 : v3(clamp01(o45.y), clamp01(o67.x), clamp01(o67.y));
        rec_store(rec, e, pos, 1, nrm, distance, col, global_mparams(f.gmat));
        return -1;
    }
    return innerHit ? 1 : 0;
}
__device__ __forceinline__ bool seg_apply_reenter(const Frame& f, int e, bool innerHit) {
    const float4 X = f.st.segX[e];
    const V3 dir = f4v(f.st.segD[e]);
    const V3 shiftedExit = f4v(X) + dir * kSegEps;
    const Hit hr = trace_dyn_inline(target_of(f.mesh[NRB_SLOT_OUTER]), TRACE_FORWARD, shiftedExit, dir);
    const bool re = hr.t < 1e30f;
    if (re || innerHit) {
        const float rd = re ? hr.t : 0.0f;
        const V3 ne = shiftedExit + dir * rd;
        f.st.segE[e] = make_float4(ne.x, ne.y, ne.z, f.st.segE[e].w + X.w + rd + 2.0f * kSegEps);
        return true;
    }
    return false;
}
__device__ __forceinline__ bool seg_apply_one(const Frame& f, int e, uint4 raw, const SurfRec& rec, unsigned int& nTrace) {
    const int st = seg_apply_decode(f, e, raw, rec);
    if (st < 0) return false;
    ++nTrace;
    return seg_apply_reenter(f, e, st != 0);
}

// One iteration, second half: decode the network output, record hits, re-enter the shell for the
// rays that go on, and compact them into the next queue.
//
// Only the rays that leave their segment without a hit trace here (10 of 32 lanes with the trace inside the
// per-ray branch, profiles/r1_v12_mid_kernels_ncu.txt): a warp decodes 32 rays per trip, parks those that must
// look for a re-entry in a 64-entry list in shared memory, and traces 32 of them at a time with every lane busy.
__device__ __forceinline__ void seg_apply_body(const Frame& f, int qin, int qout, int recIdx) {
    __shared__ int parked[kQueueThreads / 32][64];  // e | innerHit << 31
    const int n = f.st.counts[qin];
    const SurfRec& rec = f.st.rec[recIdx];
    unsigned int nTrace = 0;
    const int first = queue_start(n, 32);
    const int lane = threadIdx.x & 31;
    int* mine = parked[threadIdx.x >> 5];
    int nParked = 0;  // warp-uniform
    auto drain = [&](int take) {  // the last `take` parked rays, one per lane
        nParked -= take;
        const bool on = lane < take;
        bool again = false;
        int e = 0;
        if (on) {
            const int v = mine[nParked + lane];
            e = v & 0x7fffffff;
            again = seg_apply_reenter(f, e, v < 0);
            ++nTrace;
        }
        queue_push(f.st.queue[qout], f.st.counts + qout, again, e);
        __syncwarp();
    };
    for (int base = 0; base < n; base += gridDim.x * blockDim.x) {  // the same trip count for every thread
        const int i = base + first;
        int st = -1, e = 0;
        if (i < n) {
            e = f.st.queue[qin][i];
            st = seg_apply_decode(f, e, *reinterpret_cast<const uint4*>(f.st.netOut + static_cast<size_t>(i) * 16), rec);
        }
        const unsigned votes = __ballot_sync(0xffffffffu, st >= 0);
        if (st >= 0) mine[nParked + __popc(votes & ((1u << lane) - 1))] = e | (st << 31);
        nParked += __popc(votes);
        __syncwarp();
        if (nParked >= 32) drain(32);
    }
    if (nParked > 0) drain(nParked);
    stat_flush(f.st.stats + 0, nTrace);
}
__global__ void __launch_bounds__(kQueueThreads) k_seg_apply(const __grid_constant__ Frame f, int qin, int qout,
                                                             int recIdx) {
    seg_apply_body(f, qin, qout, recIdx);
}

// ---- the tail of the segment loop as a device-driven WHILE graph ----------------------------------
// Iterations 1..9 of traceNeuralSegmentsForRays (:2003-2081) only see rays that left a shell segment
// without a hit and re-entered the shell: usually none.  The reference decides on the host (a
// blocking count read-back per iteration); launching all nine iterations blindly costs ~36 empty
// launches per pass.  Here a CUDA conditional WHILE node repeats {exits, query, apply, control}
// while the device-resident count is non-zero.  The tail kernels read the frame block from device
// memory (it is re-uploaded per frame) so that the instantiated graph outlives the frame.
struct SegCtl {
    int qin, qout, it, curCount, recIdx;
};
__global__ void k_seg_tail_begin(SegCtl* ctl, int* counts, cudaGraphConditionalHandle h, int recIdx) {
    // after iteration 0: queue 2 -> queue 3
    const int c = counts[3];
    ctl->qin = 3; ctl->qout = 2; ctl->it = 1; ctl->curCount = c; ctl->recIdx = recIdx;
    counts[2] = 0;
    cudaGraphSetConditional(h, (c > 0 && 1 < kMaxSegIters) ? 1u : 0u);
}
__global__ void k_seg_tail_next(SegCtl* ctl, int* counts, cudaGraphConditionalHandle h) {
    const int qin = ctl->qout, qout = ctl->qin, it = ctl->it + 1;
    const int c = counts[qin];
    ctl->qin = qin; ctl->qout = qout; ctl->it = it; ctl->curCount = c;
    counts[qout] = 0;
    cudaGraphSetConditional(h, (c > 0 && it < kMaxSegIters) ? 1u : 0u);
}
__global__ void __launch_bounds__(kQueueThreads) k_seg_exits_tail(const Frame* __restrict__ f, const SegCtl* __restrict__ ctl) {
    seg_exits_body(*f, ctl->qin);
}
__global__ void __launch_bounds__(kQueueThreads) k_seg_apply_tail(const Frame* __restrict__ f, const SegCtl* __restrict__ ctl) {
    seg_apply_body(*f, ctl->qin, ctl->qout, ctl->recIdx);
}

__device__ __forceinline__ void select_closer(const SurfRec& rec, int e, bool addHit, V3 origin, V3 pos, V3 nrm, V3 alb,
                                              V3 mp, bool& flag) {
    if (!addHit) return;
    const float4 P = rec.P[e];
    const bool shellHit = __float_as_int(P.w) != 0;
    bool take = !shellHit;
    if (shellHit) take = length(pos - origin) < length(f4v(P) - origin);
    if (take) {
        const float dist = rec.N[e].w;  // the reference's select kernels leave hitDistances untouched
        rec_store(rec, e, pos, 1, nrm, dist, alb, mp);
    }
    flag = true;
}

// After the primary segment loop: hybrid mesh, closest-hit merge, path-state init, path queue.
__global__ void __launch_bounds__(kTileThreads) k_primary_finish(const __grid_constant__ Frame f) {
    const FrameParams& fp = f.fp;
    const ElemId id = tile_elem(fp);
    bool active = false;
    unsigned int nTrace = 0;
    if (id.valid) {
        uint32_t rng = rng_seed(id.pixel, fp.sampleOffset, id.s);
        const V3 dir = primary_direction(fp, id.x, id.gy, rng);
        bool hit = __float_as_int(f.st.rec[0].P[id.e].w) != 0;
        if (f.additionalEnabled && f.mesh[NRB_SLOT_ADDITIONAL].numTriangles > 0) {
            V3 pos, nrm, alb, mp;
            const bool addHit = hit_record<TRACE_FORWARD>(f, NRB_SLOT_ADDITIONAL, fp.camPos, dir, pos, nrm, alb, mp);
            ++nTrace;
            select_closer(f.st.rec[0], id.e, addHit, fp.camPos, pos, nrm, alb, mp, hit);
        }
        V3 L = v3s(0.0f);
        if (!hit) L = fp.useDirectEnvColor ? fp.directEnvColor : env_clamped(f.env, dir, fp.maxRadiance);
        f.st.T[id.e] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
        f.st.L[id.e] = make_float4(L.x, L.y, L.z, 0.0f);
        active = hit && fp.maxBounces > 0;
    }
    queue_push(f.st.queue[0], f.st.counts + 0, active, id.e);
    stat_flush(f.st.stats + 0, nTrace);
}

// Bounce, first half: sample the BRDF at every live vertex, early-terminate paths whose bounce ray
// reaches the inner shell before the outer one, find the shell entry of the survivors.
// qin -> qray (all surviving rays) and -> segment queue 2 (those that enter the shell).
//
// checkBounceEarlyTerminationKernel + traceOuterShellEntryFromRaysKernel need up to five closest-hit
// queries per vertex: ANY / FORWARD / BACKWARD on the outer shell (one fused pass, MultiTrace), ANY on
// the inner shell, FORWARD on the outer shell again from beyond the exit.  Bounce rays are incoherent
// and their traversal lengths are heavy-tailed: with one vertex per lane for the whole kernel, ncu
// showed 7.5 of 32 lanes active in the traversal loops (profiles/r1_v5_*).  This kernel is therefore a
// PERSISTENT warp scheduler: every lane runs a small state machine (fetch vertex -> sample -> trace
// -> trace -> ... -> write), the warp alternates between a convergent traversal phase and a
// convergent transition phase, and lanes whose work is done fetch the next vertex from a global
// cursor as soon as fewer than kRefillBelow lanes are still tracing.
constexpr int kRefillBelow = 22;  // 28 measured slower after the trace pruning too (0.543 -> 0.600 ms, v18)

// sampleBounceDirectionsKernel (:357-499) on the live-path queue: fully convergent apart from the
// three-way lobe choice.  A path the reference deactivates here is marked by pdf = -1.
__global__ void __launch_bounds__(kQueueThreads, 8) k_bounce_sample(const __grid_constant__ Frame f, int bounce, int qin,
                                                                    int* hostFeedback) {
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[qin];
    if (hostFeedback && blockIdx.x == 0 && threadIdx.x == 0) {  // how many paths enter this bounce: the host's hint for the next frames
        hostFeedback[0] = n;
        __threadfence_system();
    }
    const SurfRec& cur = f.st.rec[bounce == 1 ? 0 : 1];
    for (int i = queue_start(n); i < n; i += gridDim.x * blockDim.x) {
        const int e = f.st.queue[qin][i];
        const ElemId id = queue_elem(fp, e);
        const float4 P = cur.P[e], N = cur.N[e], C = cur.C[e];
        const float2 M = cur.M[e];
        const V3 prev = bounce == 1 ? v3s(0.0f) : f4v(f.st.prevD[e]);
        V3 org = v3s(0.0f), wi = v3(0.0f, 0.0f, 1.0f), weight = v3s(0.0f);
        float pdf = 0.0f;
        const bool ok = sample_vertex(f, id, f4v(P), f4v(N), f4v(C), v3(C.w, M.x, M.y), bounce == 1, prev, org, wi, weight, pdf);
        f.st.rayO[e] = make_float4(org.x, org.y, org.z, ok ? pdf : -1.0f);
        f.st.rayD[e] = make_float4(wi.x, wi.y, wi.z, N.w);
        f.st.rayW[e] = make_float4(weight.x, weight.y, weight.z, 0.0f);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(f.st.stats + 2, static_cast<unsigned long long>(n));
        // the counters this bounce fills (ray queue 1, segment queues 2 and 3, the two work cursors): their last
        // readers were earlier launches and their first writers are later ones, so no memset nodes are needed
        f.st.counts[1] = f.st.counts[2] = f.st.counts[3] = 0;
        f.st.counts[8] = f.st.counts[9] = 0;
    }
}

// Ten resident CTAs per SM (48 registers) measured best: 8 / 10 / 12 / 16 -> 0.808 / 0.791 / 0.800 / 0.921 ms (v7);
// 10 / 12 -> 0.543 / 0.572 ms with the decided traces skipped (v16).
constexpr int kBounceCtasPerSm = 10;
__global__ void __launch_bounds__(kQueueThreads, kBounceCtasPerSm) k_bounce_begin_vote(const __grid_constant__ Frame f, int bounce, int qin,
                                                                int qray) {
    const int n = f.st.counts[qin];
    int* cursor = f.st.counts + 8;
    const TraceTarget tOuter = target_of(f.mesh[NRB_SLOT_OUTER]), tInner = target_of(f.mesh[NRB_SLOT_INNER]);
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    unsigned int nTrace = 0;

    // lane state: stage 0 = no vertex; 1 = fused outer trace; 2 = inner ANY; 3 = re-entry FORWARD
    int stage = 0, e = 0;
    V3 org = v3s(0.0f), wi = v3(0.0f, 0.0f, 1.0f), entryOrigin = v3s(0.0f);
    float carried = 0.0f, tAny = 1e30f, tFwd = 1e30f, tBwd = 1e30f, baseOffset = 0.0f;
    MultiTrace tr;
    tr.ref = kTraceDone;
    const bool literal = f.literalBounceTraces != 0;
    bool more = true;  // warp-uniform: the cursor has not run past the queue
    const int warpsTotal = gridDim.x * (kQueueThreads / 32);
    const int cap = max(1, min(32, (n + warpsTotal - 1) / warpsTotal));  // lanes a warp keeps busy

    for (;;) {
        // ================= transition phase (convergent) =================
        for (;;) {
            bool finish = false, alive = false;
            float entryT = 1e30f;
            if (stage != 0 && tr.done()) {
                int next;  // 2 = inner ANY, 3 = re-entry, 5 = write
                // Traces whose outcome is decided are not run (they still count as traced): the FORWARD re-entry
                // trace starts beyond the BACKWARD exit of a ray on which the fused trace found no front face at
                // all - its candidates are a subset of that trace's, so it finds nothing either; and the inner-shell
                // trace only matters where it is nearer than the outer-shell hit.  NRB_OPT_LITERAL_BOUNCE_TRACES
                // runs them all, as the reference does.
                if (stage == 1) {
                    tAny = tr.best0; tFwd = tr.best1;
                    // outside the literal setting the BACKWARD class is not traced: its distance is only asked "is there
                    // one" when there is no front face, and then the nearest face of any kind is a back face
                    tBwd = literal ? tr.best2 : (tFwd < 1e30f ? 1e30f : tAny);
                    const bool wouldReenter = !(tFwd < 1e30f) && tBwd < 1e30f;
                    nTrace += (carried > 0.0f ? 1u : 0u) + 1u + (tFwd >= 1e30f ? 1u : 0u);
                    const int after = tFwd < 1e30f ? 5 : (tBwd < 1e30f ? (literal ? 3 : 5) : 5);
                    next = carried > 0.0f ? 2 : after;
                    if (!literal) {
                        if (next == 2 && !(tAny < 1e30f)) { next = after; ++nTrace; }  // nothing to be nearer than
                        if (next == 5 && wouldReenter) ++nTrace;
                    }
                    entryT = tFwd;
                    alive = true;
                } else if (stage == 2) {
                    const float ht = tr.best0;
                    const int after = tFwd < 1e30f ? 5 : (tBwd < 1e30f ? (literal ? 3 : 5) : 5);
                    if (!literal && !(tFwd < 1e30f) && tBwd < 1e30f && !(ht < 1e30f && tAny < 1e30f && ht < tAny)) ++nTrace;
                    if (ht < 1e30f && tAny < 1e30f && ht < tAny) { next = 5; alive = false; }
                    else { next = after; alive = true; entryT = tFwd; }
                } else {
                    next = 5;
                    alive = true;
                    entryT = tr.best1;
                }
                if (next == 2) {
                    tr.begin(tInner, 1, org, wi, literal ? 1e30f : tAny);
                    ++nTrace;
                    stage = 2;
                } else if (next == 3) {
                    baseOffset = tBwd + kSegEps;
                    entryOrigin = org + wi * baseOffset;
                    tr.begin(tOuter, 2, entryOrigin, wi);
                    ++nTrace;
                    stage = 3;
                } else {
                    finish = true;
                    stage = 0;
                }
            }
            bool enters = false;
            if (finish && alive) {
                rec_clear(f.st.rec[1], e);
                if (entryT < 1e30f) {
                    const V3 p = entryOrigin + wi * entryT;
                    f.st.segE[e] = make_float4(p.x, p.y, p.z, baseOffset + entryT);
                    f.st.segD[e] = make_float4(wi.x, wi.y, wi.z, 0.0f);
                    enters = true;
                }
            }
            queue_push(f.st.queue[qray], f.st.counts + qray, finish && alive, e);
            queue_push(f.st.queue[2], f.st.counts + 2, enters, e);

            // idle lanes fetch the next vertices, up to `cap` busy lanes per warp.  A launch with fewer vertices
            // than lanes (later bounces: a few thousand survivors) is latency-bound: a warp that holds r rays needs
            // about twice the steps of its longest ray in vote iterations whatever r is, so the vertices are spread
            // one per warp instead of 32 to the first warps (measured: 131 -> 4x us for 3.3 K vertices).
            const unsigned idle = __ballot_sync(full, stage == 0);
            const int take = more ? min(__popc(idle), cap - (32 - __popc(idle))) : 0;
            if (take > 0) {
                int base = 0;
                const int leader = __ffs(idle) - 1;
                if (lane == leader) base = atomicAdd(cursor, take);
                base = __shfl_sync(full, base, leader);
                more = base + take < n;
                const int rank = __popc(idle & ((1u << lane) - 1));
                const int i = base + rank;
                if (stage == 0 && rank < take && i < n) {
                    e = f.st.queue[qin][i];
                    const float4 O = f.st.rayO[e], D = f.st.rayD[e];  // written by k_bounce_sample
                    org = f4v(O);
                    wi = f4v(D);
                    carried = D.w;
                    if (O.w > 0.0f) {
                        baseOffset = 0.0f;
                        entryOrigin = org;
                        tr.begin(tOuter, literal ? ((carried > 0.0f ? 1 : 0) | 2 | 4) : (1 | 2), org, wi);
                        stage = 1;
                    }
                }
            }
            // again while a lane still has a finished trace to act on, or the warp is below its share with work left
            const unsigned busyB = __ballot_sync(full, stage != 0);
            const bool pending = (stage != 0 && tr.done()) || (more && take > 0 && __popc(busyB) < cap && busyB != full);
            if (!__any_sync(full, pending)) break;
        }
        if (!__any_sync(full, stage != 0)) break;  // nothing in flight and nothing left to fetch

        // ================= traversal phase (convergent) =================
        // Every iteration the warp votes: the unit kind wanted by most lanes runs, the others wait (and
        // pile up until they are the majority), so each executed instruction carries a large lane group
        // instead of node steps, triangle tests and pops serialising at a few lanes each.
        for (;;) {
            const int k = tr.kind();
            const unsigned bN = __ballot_sync(full, k == 0), bT = __ballot_sync(full, k == 1), bP = __ballot_sync(full, k == 2);
            const int busy = __popc(bN | bT | bP);
            if (busy == 0 || (more && busy < min(kRefillBelow, cap))) break;
            if (busy == 1) {
                // a lone ray (short launches: one per warp; the end of the drain): nobody to vote with, and the
                // latency of this warp's instruction stream is all there is - run the ray to its end directly
                if (k < 3) {
                    do {
                        const int kk = tr.kind();
                        if (kk == 0) tr.step_node();
                        else if (kk == 1) tr.step_tri();
                        else tr.step_pop();
                    } while (!tr.done());
                }
                break;
            }
            const int nN = __popc(bN), nT = __popc(bT), nP = __popc(bP);
            // Units come in runs (a descent is followed by a descent, a leaf block by the next block of the
            // leaf), so a lane that stays in the voted kind does a second unit before the next vote: units per
            // vote 1 / 2 / 3 / 4 for nodes -> 0.760 / 0.734 / 0.739 / 0.739 ms for this kernel at 1080p, then
            // 2 / 3 / 4 for triangle blocks -> 0.698 / 0.696 / 0.726 ms; chaining a pop after the last block
            // or two pops per vote measured slower (0.715 / 0.698 ms).
            if (nP >= 8 || (nP >= nN && nP >= nT)) {
                if (k == 2) tr.step_pop();
            } else if (nN >= nT) {
                if (k == 0) {
                    tr.step_node();
                    if (tr.kind() == 0) tr.step_node();
                }
            } else {
                if (k == 1) {
                    tr.step_tri();
                    if (tr.kind() == 1) tr.step_tri();
                }
            }
        }
    }
    stat_flush(f.st.stats + 0, nTrace);
}

// The bounce entry since v26 (NRB_OPT_BOUNCE_ENTRY = 0): separate, PRUNED traces instead of the fused three-class walk of
// k_bounce_begin_vote, whose FORWARD class finds nothing on 96 % of the rays and so never prunes.  A warp takes 32
// vertices from the cursor.  (1) The outer shell's nearest face of any kind - a closest-hit walk that stops at the exit
// a few percent of the object away.  (2) The nearest FRONT face: if the nearest face is a front face, that is it; if
// it is a back face, no front face lies nearer (it would have been the nearest face), so the FORWARD walk skips every
// subtree the ray leaves before that distance (trace_sync's tMin) - which is the whole neighbourhood of the origin -
// and sees only what lies beyond the exit.  (3) The inner shell as far as the outer hit.  Each is a trace_sync walk
// (reconverged every few steps).  Same decisions and counters as the voting kernel (tests compare the two):
// 0.452 -> 0.342 ms at 1080p.  Cost split measured by leaving traces out: (1) 0.205, (3) 0.09, (2) 0.13 unbounded,
// ~0.045 with the lower bound.
constexpr int kBounceEntryCtasPerSm = 8;  // 6 / 8 / 10: 0.456 / 0.424 / 0.424 ms before the lower bound; 8 / 10 / 12: 0.341 / 0.343 / 0.348 with it
__global__ void __launch_bounds__(kQueueThreads, kBounceEntryCtasPerSm) k_bounce_begin(const __grid_constant__ Frame f, int bounce,
                                                                                      int qin, int qray) {
    const int n = f.st.counts[qin];
    int* cursor = f.st.counts + 8;
    const TraceTarget tOuter = target_of(f.mesh[NRB_SLOT_OUTER]), tInner = target_of(f.mesh[NRB_SLOT_INNER]);
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const bool literal = f.literalBounceTraces != 0;
    unsigned int nTrace = 0;
    for (;;) {
        int base = 0;
        if (lane == 0) base = atomicAdd(cursor, 32);
        base = __shfl_sync(full, base, 0);
        if (base >= n) break;
        const int i = base + lane;
        int e = 0;
        V3 org = v3s(0.0f), wi = v3(0.0f, 0.0f, 1.0f);
        float carried = 0.0f;
        bool has = false;
        if (i < n) {
            e = f.st.queue[qin][i];
            const float4 O = f.st.rayO[e], D = f.st.rayD[e];
            org = f4v(O);
            wi = f4v(D);
            carried = D.w;
            has = O.w > 0.0f;
        }
        V3 entryOrigin = org;
        float baseOffset = 0.0f;
        // nearest face of any kind; if it is a front face it is also the nearest front face, and if it is a back face
        // no front face lies nearer: the FORWARD search then skips everything the ray leaves before that distance
        const Hit hAny = trace_sync(tOuter, TRACE_ANY, has, org, wi);
        const float tAny = hAny.t;
        bool anyIsFront = false;
        if (has && tAny < 1e30f) {
            const TriBlock& tb = tOuter.tris[hAny.prim];
            anyIsFront = dot(v3(__ldg(&tb.p.w), __ldg(&tb.q.w), __ldg(&tb.r.w)), wi) < 0.0f;
        }
        const bool needFwd = has && tAny < 1e30f && !anyIsFront;
        float tFwd = trace_sync<false, true>(tOuter, TRACE_FORWARD, needFwd, org, wi, 1e30f, nullptr, tAny).t;
        if (!needFwd) tFwd = anyIsFront ? tAny : 1e30f;
        float tBwd = tFwd < 1e30f ? 1e30f : tAny;
        if (literal) tBwd = trace_sync(tOuter, TRACE_BACKWARD, has, org, wi).t;
        const int after = tFwd < 1e30f ? 5 : (tBwd < 1e30f ? (literal ? 3 : 5) : 5);
        int next = 5;
        bool alive = has;
        float entryT = tFwd;
        if (has) {
            const bool wouldReenter = !(tFwd < 1e30f) && tBwd < 1e30f;
            nTrace += (carried > 0.0f ? 1u : 0u) + 1u + (tFwd >= 1e30f ? 1u : 0u);
            next = carried > 0.0f ? 2 : after;
            if (!literal) {
                if (next == 2 && !(tAny < 1e30f)) { next = after; ++nTrace; }
                if (next == 5 && wouldReenter) ++nTrace;
            }
        }
        const bool need2 = has && next == 2;
        if (need2) ++nTrace;
        const float ht = trace_sync(tInner, TRACE_ANY, need2, org, wi, literal ? 1e30f : tAny).t;
        if (need2) {
            const bool ends = ht < 1e30f && tAny < 1e30f && ht < tAny;
            if (!literal && !(tFwd < 1e30f) && tBwd < 1e30f && !ends) ++nTrace;
            if (ends) { next = 5; alive = false; }
            else { next = after; entryT = tFwd; }
        }
        if (__any_sync(full, has && next == 3)) {  // literal setting only: the re-entry trace from beyond the exit
            const bool need3 = has && next == 3;
            if (need3) {
                baseOffset = tBwd + kSegEps;
                entryOrigin = org + wi * baseOffset;
                ++nTrace;
            }
            const float t3 = trace_sync(tOuter, TRACE_FORWARD, need3, entryOrigin, wi).t;
            if (need3) entryT = t3;
        }
        bool enters = false;
        if (alive) {
            rec_clear(f.st.rec[1], e);
            if (entryT < 1e30f) {
                const V3 p = entryOrigin + wi * entryT;
                f.st.segE[e] = make_float4(p.x, p.y, p.z, baseOffset + entryT);
                f.st.segD[e] = make_float4(wi.x, wi.y, wi.z, 0.0f);
                enters = true;
            }
        }
        queue_push(f.st.queue[qray], f.st.counts + qray, alive, e);
        queue_push(f.st.queue[2], f.st.counts + 2, enters, e);
    }
    stat_flush(f.st.stats + 0, nTrace);
}

// Bounce, second half: hybrid mesh + merge, throughput/radiance update, Russian roulette.
__global__ void __launch_bounds__(kQueueThreads) k_bounce_finish(const __grid_constant__ Frame f, int bounce, int qray,
                                                                 int qout) {
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[qray];
    const SurfRec& rec = f.st.rec[1];
    unsigned int nTrace = 0;
    const bool useAdd = f.additionalEnabled && f.mesh[NRB_SLOT_ADDITIONAL].numTriangles > 0;
    const int first = queue_start(n);
    for (int base = 0; base < n; base += gridDim.x * blockDim.x) {  // the same trip count for every thread
        const int i = base + first;
        bool alive = false;
        int e = 0;
        if (i < n) {
            e = f.st.queue[qray][i];
            const ElemId id = queue_elem(fp, e);
            const V3 org = f4v(f.st.rayO[e]), wi = f4v(f.st.rayD[e]), weight = f4v(f.st.rayW[e]);
            bool hit = __float_as_int(rec.P[e].w) != 0;
            if (useAdd) {
                V3 pos, nrm, alb, mp;
                const bool addHit = hit_record<TRACE_ANY>(f, NRB_SLOT_ADDITIONAL, org, wi, pos, nrm, alb, mp);
                ++nTrace;
                select_closer(rec, e, addHit, org, pos, nrm, alb, mp, hit);
            }
            V3 T = f4v(f.st.T[e]), L = v3s(0.0f);
            bool touchedL;
            if (!hit) L = f4v(f.st.L[e]);
            alive = integrate_vertex(f, id, bounce, hit, wi, weight, T, L, touchedL);
            if (touchedL) f.st.L[e] = make_float4(L.x, L.y, L.z, 0.0f);
            if (alive) {
                f.st.T[e] = make_float4(T.x, T.y, T.z, 0.0f);
                f.st.prevD[e] = make_float4(wi.x, wi.y, wi.z, 0.0f);
            }
        }
        queue_push(f.st.queue[qout], f.st.counts + qout, alive, e);
    }
    stat_flush(f.st.stats + 0, nTrace);
}

// ---------------------------------------------------------------------------------------------

// The segment loop of traceNeuralSegmentsForRays.  Iteration 0 is launched directly (full grids,
// frame block in the constant bank).  The remaining iterations run either as the WHILE graph above
// (tail != nullptr) or, for the per-launch profiler and as the fallback, as nine fixed one-wave
// launches whose grid-stride loops make any count correct.
static int seg_iteration(const Frame& f, const NetDev& net, int recIdx, int qin, int grid, cudaStream_t s, bool first) {
    const int qout = qin == 2 ? 3 : 2;
    // iteration 0 of a pass finds queue 3 empty (frame start / k_bounce_sample); the fixed loop resets per iteration
    if (!first) cudaMemsetAsync(f.st.counts + qout, 0, sizeof(int), s);
    k_seg_exits<<<grid, kQueueThreads, 0, s>>>(f, qin, qin == 2 && grid > sm_count() ? 1 : 0);  // iteration 0 of a pass
    NRB_LAUNCHED("k_seg_exits", s);
    int launches = 1 + launch_neural_query(net, f.st.netIn, static_cast<size_t>(f.fp.nElems), f.st.counts + qin, f.st.netOut, s);
    k_seg_apply<<<grid, kQueueThreads, 0, s>>>(f, qin, qout, recIdx);
    NRB_LAUNCHED("k_seg_apply", s);
    return launches + 1;
}

void SegTail::release() {
    for (int i = 0; i < 2; ++i) {
        if (exec[i]) cudaGraphExecDestroy(exec[i]);
        if (graph[i]) cudaGraphDestroy(graph[i]);
        exec[i] = nullptr;
        graph[i] = nullptr;
    }
    if (capture) cudaStreamDestroy(capture);
    capture = nullptr;
    if (dFrame) cudaFree(dFrame);
    if (dCtl) cudaFree(dCtl);
    dFrame = dCtl = nullptr;
    keyBytes = 0;
}

// Build graph[recIdx]: k_seg_tail_begin -> WHILE { exits, query, apply, k_seg_tail_next }.
static bool build_tail(SegTail& t, const Frame& f, const NetDev& net, int recIdx) {
    cudaGraph_t g = nullptr;
    if (cudaGraphCreate(&g, 0) != cudaSuccess) return false;
    t.graph[recIdx] = g;
    cudaGraphConditionalHandle h;
    if (cudaGraphConditionalHandleCreate(&h, g, 0, cudaGraphCondAssignDefault) != cudaSuccess) return false;
    SegCtl* ctl = static_cast<SegCtl*>(t.dCtl) + recIdx;
    int* counts = f.st.counts;
    cudaGraphNode_t begin = nullptr;
    {
        void* args[] = {&ctl, &counts, &h, &recIdx};
        cudaKernelNodeParams kp{};
        kp.func = reinterpret_cast<void*>(k_seg_tail_begin);
        kp.gridDim = dim3(1);
        kp.blockDim = dim3(1);
        kp.kernelParams = args;
        if (cudaGraphAddKernelNode(&begin, g, nullptr, 0, &kp) != cudaSuccess) return false;
    }
    cudaGraphNodeParams cp{};
    cp.type = cudaGraphNodeTypeConditional;
    cp.conditional.handle = h;
    cp.conditional.type = cudaGraphCondTypeWhile;
    cp.conditional.size = 1;
    cudaGraphNode_t loop = nullptr;
    if (cudaGraphAddNode(&loop, g, &begin, 1, &cp) != cudaSuccess) return false;
    cudaGraph_t body = cp.conditional.phGraph_out[0];
    if (cudaStreamBeginCaptureToGraph(t.capture, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed) != cudaSuccess) return false;
    const Frame* df = static_cast<const Frame*>(t.dFrame);
    k_seg_exits_tail<<<sm_count(), kQueueThreads, 0, t.capture>>>(df, ctl);
    const int q = launch_neural_query(net, f.st.netIn, static_cast<size_t>(f.fp.nElems), &ctl->curCount, f.st.netOut, t.capture);
    k_seg_apply_tail<<<sm_count(), kQueueThreads, 0, t.capture>>>(df, ctl);
    k_seg_tail_next<<<1, 1, 0, t.capture>>>(ctl, counts, h);
    cudaGraph_t done = nullptr;
    if (cudaStreamEndCapture(t.capture, &done) != cudaSuccess || q < 0) return false;
    if (cudaGraphInstantiate(&t.exec[recIdx], g, 0) != cudaSuccess) return false;
    return true;
}

// Make sure the graphs match this frame's buffers and network; (re)build them if not.
static bool prepare_tail(SegTail& t, const Frame& f, const NetDev& net, cudaStream_t s) {
    if (t.failed) return false;
    unsigned char key[512] = {};
    size_t n = 0;
    auto put = [&](const void* p, size_t bytes) {
        if (n + bytes <= sizeof(key)) std::memcpy(key + n, p, bytes);
        n += bytes;
    };
    put(&net, sizeof(NetDev));
    put(&f.st.netIn, sizeof(void*));
    put(&f.st.netOut, sizeof(void*));
    put(&f.st.counts, sizeof(void*));
    put(&f.fp.nElems, sizeof(int));
    if (n > sizeof(key)) { t.failed = true; return false; }
    if (t.exec[0] && t.exec[1] && n == t.keyBytes && std::memcmp(key, t.key, n) == 0) return true;
    StageProfiler* prof = get_stage_profiler();  // no per-launch events inside a capture
    set_stage_profiler(nullptr);
    struct Restore {
        StageProfiler* p;
        ~Restore() { set_stage_profiler(p); }
    } restore{prof};
    cudaStreamSynchronize(s);  // graphs of the previous key may still be running
    for (int i = 0; i < 2; ++i) {
        if (t.exec[i]) cudaGraphExecDestroy(t.exec[i]);
        if (t.graph[i]) cudaGraphDestroy(t.graph[i]);
        t.exec[i] = nullptr;
        t.graph[i] = nullptr;
    }
    bool ok = true;
    if (!t.capture) ok = cudaStreamCreateWithFlags(&t.capture, cudaStreamNonBlocking) == cudaSuccess;
    if (ok && !t.dFrame) ok = cudaMalloc(&t.dFrame, sizeof(Frame)) == cudaSuccess;
    if (ok && !t.dCtl) ok = cudaMalloc(&t.dCtl, 2 * sizeof(SegCtl)) == cudaSuccess;
    if (ok) {
        // one empty query outside capture: the launchers set their function attributes on first use
        cudaMemsetAsync(t.dCtl, 0, 2 * sizeof(SegCtl), s);
        launch_neural_query(net, f.st.netIn, static_cast<size_t>(f.fp.nElems), &static_cast<SegCtl*>(t.dCtl)->curCount, f.st.netOut, s);
        cudaStreamSynchronize(s);
    }
    ok = ok && build_tail(t, f, net, 0) && build_tail(t, f, net, 1);
    if (!ok) {
        cudaGetLastError();
        t.failed = true;
        fprintf(stderr, "[nrb] conditional-graph segment tail unavailable; using the fixed iteration loop\n");
        return false;
    }
    std::memcpy(t.key, key, n);
    t.keyBytes = n;
    return true;
}

// ---- the tail of the segment loop as ONE launch ------------------------------------------------------
// The rays that leave iteration 0 of a pass without a hit and re-enter the shell (queue 3) are few - usually
// none - and independent of each other, so nothing needs a grid-wide step between their iterations: a warp
// takes 16 of them and carries them through exits -> network -> apply until each has hit, left, or used up
// its iterations.  The network runs as the warp-autonomous mma.sync tile of nrb_query_warp.cuh against weights
// the CTA loads once.  With an empty queue the launch returns at once: 4 such launches per frame replace four
// conditional-graph launches of ~23 us each (profiles/: segment_tail_graph 0.094 ms per frame).
template <int WIDTH>
__global__ void __launch_bounds__(512, 1)
    k_seg_tail(const __grid_constant__ Frame f, const __grid_constant__ NetDev net, int recIdx, int ek, int warps) {
    using namespace qwarp;
    const int n = f.st.counts[3];
    // path queue 0 was read by the stages before the segment loop and is refilled by k_bounce_finish after it
    if (recIdx == 1 && blockIdx.x == 0 && threadIdx.x == 0) f.st.counts[0] = 0;
    if (n == 0 || blockIdx.x * warps * kRows >= n) return;
    extern __shared__ __align__(16) unsigned char tailSmem[];
    const int H = net.nHidden;
    const int ld0 = ek + 8, ldw = WIDTH + 8, ldA = (ek > WIDTH ? ek : WIDTH) + 8;
    __half* w0 = reinterpret_cast<__half*>(tailSmem);
    __half* wh = w0 + WIDTH * ld0;
    __half* wl = wh + (H - 1) * WIDTH * ldw;
    unsigned char* perWarp = reinterpret_cast<unsigned char*>(wl + 16 * ldw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    preload_weights<WIDTH>(net, ek, w0, wh, wl);
    __syncthreads();
    if (warp >= warps) return;
    const size_t perWarpBytes = static_cast<size_t>(kRows) * ldA * 2 + static_cast<size_t>(kRows) * 16 * 4 + kRows * 16 * 2;
    __half* act = reinterpret_cast<__half*>(perWarp + warp * perWarpBytes);
    float* inp = reinterpret_cast<float*>(act + kRows * ldA);          // [16][16] fp32 inputs
    __half* outRows = reinterpret_cast<__half*>(inp + kRows * 16);      // [16][16] fp16 outputs
    const SurfRec& rec = f.st.rec[recIdx];
    unsigned int nTrace = 0, nQuery = 0;
    const int tiles = (n + kRows - 1) / kRows;
    for (int tile = blockIdx.x * warps + warp; tile < tiles; tile += gridDim.x * warps) {
        const int i = tile * kRows + lane;
        bool alive = lane < kRows && i < n;
        const int e = alive ? f.st.queue[3][i] : 0;
        for (int it = 1; it < kMaxSegIters; ++it) {
            if (lane < kRows)
                for (int c = 0; c < 16; ++c) inp[lane * 16 + c] = 0.0f;
            if (alive) {
                seg_exit_one(f, e, inp + lane * 16);
                ++nQuery;
            }
            __syncwarp();
            query_tile<WIDTH>(net, ek, w0, wh, wl, act, inp, outRows, 16, kRows);
            if (alive) alive = seg_apply_one(f, e, *reinterpret_cast<const uint4*>(outRows + lane * 16), rec, nTrace);
            __syncwarp();
            if (!__any_sync(0xffffffffu, alive)) break;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        nTrace += __shfl_xor_sync(0xffffffffu, nTrace, o);
        nQuery += __shfl_xor_sync(0xffffffffu, nQuery, o);
    }
    if (lane == 0 && (nTrace | nQuery)) {
        atomicAdd(f.st.stats + 0, static_cast<unsigned long long>(nTrace) + 2ull * nQuery);
        atomicAdd(f.st.stats + 1, static_cast<unsigned long long>(nQuery));
    }
}

// ---- every bounce after the first as ONE launch --------------------------------------------------------
// On the bench scene bounce 1 carries 780 K path vertices, bounce 2 three thousand and bounce 3 about fifty, and
// each of those bounces still cost eight dependent launches whose time was the latency of a handful of warps
// (~0.1 ms per bounce).  Paths are independent of each other, so nothing needs a
// grid-wide step between their bounces: a warp takes up to 16 of the surviving paths - spread one per warp while
// there are fewer paths than warps - and carries them through
//   sample (sampleBounceDirections :357-499) -> early termination + shell entry (:503-553, :773-852)
//   -> segment loop (exits, network, apply, re-entry; :858-1123, :1456-1524) -> hybrid mesh + integrate (:1235-1450, :621-675)
// bounce after bounce until each has ended.  The network runs as the warp-autonomous mma.sync tile with its weights
// read straight from global memory (query_tile_gmem: a few tiles per warp do not repay a 170 KB preload per CTA, and
// without that shared-memory carve-out the L1 keeps the shells' BVH nodes - with it every node fetch of these few,
// latency-bound warps went to L2); the traversals are the exact per-mode trace_multi / trace_dyn loops.  Same results
// as the per-stage kernels up to the order of the fp32 sums inside the network (fp16 tolerance).
// The host launches it instead of the per-bounce stages when the previous frames entered bounce 2 with few paths
// (the count travels back through mapped host memory, no synchronisation); it is correct for any count.
constexpr int kPathTailWarps = 8;
template <int WIDTH>
__global__ void __launch_bounds__(kPathTailWarps * 32, 2)
    k_path_tail(const __grid_constant__ Frame f, const __grid_constant__ NetDev net, int firstBounce, int ek,
                int* hostFeedback) {
    constexpr int warps = kPathTailWarps;
    using namespace qwarp;
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[0];
    if (blockIdx.x == 0 && threadIdx.x == 0 && hostFeedback) {
        hostFeedback[0] = n;
        __threadfence_system();
    }
    const int totalWarps = gridDim.x * warps;
    int per = (n + totalWarps - 1) / totalWarps;  // paths per warp trip: one while the paths are fewer than the warps
    per = per < 1 ? 1 : (per > kRows ? kRows : per);
    if (n == 0 || blockIdx.x * warps * per >= n) return;
    extern __shared__ __align__(16) unsigned char tailSmem[];
    const int ldA = (ek > WIDTH ? ek : WIDTH) + 8;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t perWarpBytes = static_cast<size_t>(kRows) * ldA * 2 + static_cast<size_t>(kRows) * 16 * 4 + kRows * 16 * 2;
    __half* act = reinterpret_cast<__half*>(tailSmem + warp * perWarpBytes);
    float* inp = reinterpret_cast<float*>(act + kRows * ldA);
    __half* outRows = reinterpret_cast<__half*>(inp + kRows * 16);
    const TraceTarget tOuter = target_of(f.mesh[NRB_SLOT_OUTER]), tInner = target_of(f.mesh[NRB_SLOT_INNER]);
    const bool literal = f.literalBounceTraces != 0;
    const bool useAdd = f.additionalEnabled && f.mesh[NRB_SLOT_ADDITIONAL].numTriangles > 0;
    const SurfRec& rec = f.st.rec[1];
    unsigned int nTrace = 0, nQuery = 0, nVert = 0;
    const unsigned full = 0xffffffffu;
    // path j of a trip lives on lane 2j (tile row j); the odd lane beside it is its partner in the segment exits, where
    // the two traces of a segment (outer and inner shell) run side by side instead of one after the other
    const int row = lane >> 1;
    const bool even = (lane & 1) == 0;
    for (int base = (blockIdx.x * warps + warp) * per; base < n; base += totalWarps * per) {
        bool alive = even && row < per && base + row < n;
        const int e = alive ? f.st.queue[0][base + row] : 0;
        const ElemId id = queue_elem(fp, e);
        for (int b = firstBounce; b <= fp.maxBounces && __any_sync(full, alive); ++b) {
            // ---- sample + early termination + shell entry
            V3 org = v3s(0.0f), wi = v3(0.0f, 0.0f, 1.0f), weight = v3s(0.0f);
            bool enters = false;
            if (alive) {
                const SurfRec& cur = f.st.rec[b == 1 ? 0 : 1];
                const float4 P = cur.P[e], N = cur.N[e], C = cur.C[e];
                const float2 M = cur.M[e];
                const V3 prev = b == 1 ? v3s(0.0f) : f4v(f.st.prevD[e]);
                float pdf = 0.0f;
                ++nVert;
                alive = sample_vertex(f, id, f4v(P), f4v(N), f4v(C), v3(C.w, M.x, M.y), b == 1, prev, org, wi, weight, pdf);
                if (alive) {
                    const float carried = N.w;
                    float tAny, tFwd, tBwd;
                    trace_multi(tOuter, (carried > 0.0f ? 1 : 0) | 2 | 4, org, wi, tAny, tFwd, tBwd);
                    nTrace += (carried > 0.0f ? 2u : 0u) + 1u + (tFwd >= 1e30f ? 1u : 0u);
                    if (carried > 0.0f && (literal || tAny < 1e30f)) {
                        const Hit hi = trace_dyn(tInner, TRACE_ANY, org, wi);
                        if (hi.t < 1e30f && tAny < 1e30f && hi.t < tAny) alive = false;
                    }
                    float entryT = tFwd, baseOffset = 0.0f;
                    V3 entryOrigin = org;
                    if (alive && !(tFwd < 1e30f) && tBwd < 1e30f) {
                        ++nTrace;
                        if (literal) {  // restart beyond the exit and look for a front face again (:806-838)
                            baseOffset = tBwd + kSegEps;
                            entryOrigin = org + wi * baseOffset;
                            entryT = trace_dyn(tOuter, TRACE_FORWARD, entryOrigin, wi).t;
                        }
                    }
                    if (alive) {
                        rec_clear(rec, e);
                        if (entryT < 1e30f) {
                            const V3 p = entryOrigin + wi * entryT;
                            f.st.segE[e] = make_float4(p.x, p.y, p.z, baseOffset + entryT);
                            f.st.segD[e] = make_float4(wi.x, wi.y, wi.z, 0.0f);
                            enters = true;
                        }
                    }
                }
            }
            // ---- segment loop of the rays that entered the shell
            for (int it = 0; it < kMaxSegIters && __any_sync(full, enters); ++it) {
                if (lane < kRows)
                    for (int c = 0; c < 16; ++c) inp[lane * 16 + c] = 0.0f;
                __syncwarp();
                {
                    const int eP = __shfl_sync(full, e, lane & ~1);
                    const bool entP = __shfl_sync(full, enters ? 1 : 0, lane & ~1) != 0;
                    seg_exit_one(f, eP, inp + row * 16, lane & 1, entP);
                    if (enters) ++nQuery;
                }
                __syncwarp();
                query_tile_gmem<WIDTH>(net, ek, act, inp, outRows, 16, kRows, per);
                if (enters) enters = seg_apply_one(f, e, *reinterpret_cast<const uint4*>(outRows + row * 16), rec, nTrace);
                __syncwarp();
            }
            // ---- hybrid mesh, throughput / radiance, Russian roulette
            if (alive) {
                bool hit = __float_as_int(rec.P[e].w) != 0;
                if (useAdd) {
                    V3 pos, nrm, alb, mp;
                    const bool addHit = hit_record<TRACE_ANY>(f, NRB_SLOT_ADDITIONAL, org, wi, pos, nrm, alb, mp);
                    ++nTrace;
                    select_closer(rec, e, addHit, org, pos, nrm, alb, mp, hit);
                }
                V3 T = f4v(f.st.T[e]), L = v3s(0.0f);
                bool touchedL;
                if (!hit) L = f4v(f.st.L[e]);
                alive = integrate_vertex(f, id, b, hit, wi, weight, T, L, touchedL);
                if (touchedL) f.st.L[e] = make_float4(L.x, L.y, L.z, 0.0f);
                if (alive) {
                    f.st.T[e] = make_float4(T.x, T.y, T.z, 0.0f);
                    f.st.prevD[e] = make_float4(wi.x, wi.y, wi.z, 0.0f);
                }
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        nTrace += __shfl_xor_sync(full, nTrace, o);
        nQuery += __shfl_xor_sync(full, nQuery, o);
        nVert += __shfl_xor_sync(full, nVert, o);
    }
    if (lane == 0 && (nTrace | nQuery | nVert)) {
        atomicAdd(f.st.stats + 0, static_cast<unsigned long long>(nTrace) + 2ull * nQuery);
        atomicAdd(f.st.stats + 1, static_cast<unsigned long long>(nQuery));
        atomicAdd(f.st.stats + 2, static_cast<unsigned long long>(nVert));
    }
}

// Shared-memory plan of k_seg_tail (weights + per-warp tiles); 0 warps = shape not supported by the mma.sync tile.
static size_t seg_tail_smem(const NetDev& net, int* warpsOut, int* ekOut) {
    int w = 0;
    *warpsOut = 0;
    if (neural_query_smem(net, &w) == 0 || w <= 0) return 0;
    const qwarp::SmemPlan p = qwarp::plan_smem(net);
    const size_t wbytes = p.weightHalfs * 2, per = p.perWarpBytes + qwarp::kRows * 16 * 2;
    int warps = static_cast<int>((227 * 1024 - wbytes) / per);
    if (warps > 16) warps = 16;
    if (warps <= 0) return 0;
    *warpsOut = warps;
    *ekOut = p.ek;
    return wbytes + per * warps;
}

void init_frame_kernels() {
    cudaFuncSetAttribute(k_seg_tail<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    cudaFuncSetAttribute(k_seg_tail<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    cudaFuncSetAttribute(k_path_tail<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cudaFuncSetAttribute(k_path_tail<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
}

// Bounces firstBounce..maxBounces in one launch (see k_path_tail).
static bool path_tail_plan(const NetDev& net, int* ekOut, size_t* smemOut) {
    int warps = 0, ek = 0;
    if (seg_tail_smem(net, &warps, &ek) == 0) return false;  // shapes the mma.sync tile supports
    const int ldA = (ek > net.width ? ek : net.width) + 8;
    const size_t smem = static_cast<size_t>(kPathTailWarps) * (qwarp::kRows * ldA * 2 + qwarp::kRows * 16 * 4 + qwarp::kRows * 16 * 2);
    if (smem > 64 * 1024) return false;
    *ekOut = ek;
    *smemOut = smem;
    return true;
}

static bool launch_path_tail(const Frame& f, const NetDev& net, int firstBounce, int* hostFeedback, cudaStream_t s) {
    int ek = 0;
    size_t smem = 0;
    if (!path_tail_plan(net, &ek, &smem)) return false;
    const int grid = sm_count() * 2;
    if (net.width == 128) k_path_tail<128><<<grid, kPathTailWarps * 32, smem, s>>>(f, net, firstBounce, ek, hostFeedback);
    else k_path_tail<64><<<grid, kPathTailWarps * 32, smem, s>>>(f, net, firstBounce, ek, hostFeedback);
    NRB_LAUNCHED("k_path_tail", s);
    return true;
}

static bool launch_seg_tail(const Frame& f, const NetDev& net, int recIdx, cudaStream_t s) {
    int warps = 0, ek = 0;
    const size_t smem = seg_tail_smem(net, &warps, &ek);
    if (smem == 0) return false;
    if (net.width == 128) k_seg_tail<128><<<sm_count(), 512, smem, s>>>(f, net, recIdx, ek, warps);
    else k_seg_tail<64><<<sm_count(), 512, smem, s>>>(f, net, recIdx, ek, warps);
    NRB_LAUNCHED("k_seg_tail", s);
    return true;
}

static int run_segments(const Frame& f, const NetDev& net, int recIdx, cudaStream_t s, SegTail* tail, bool fused,
                        bool* fusedRan = nullptr) {
    int launches = seg_iteration(f, net, recIdx, 2, queue_grid(f.fp.nElems), s, true);
    const bool ran = fused && launch_seg_tail(f, net, recIdx, s);
    if (fusedRan) *fusedRan = ran;
    if (ran) return launches + 1;
    if (tail) {
        cudaGraphLaunch(tail->exec[recIdx], s);
        NRB_LAUNCHED("segment_tail_graph", s);
        return launches + 1;
    }
    int qin = 3;
    for (int it = 1; it < kMaxSegIters; ++it) {
        launches += seg_iteration(f, net, recIdx, qin, sm_count(), s, false);
        qin = qin == 2 ? 3 : 2;
    }
    return launches;
}

int launch_neural_frame(const Frame& f, const NetDev& net, bool lambert, cudaStream_t s, SegTail* tail, bool fusedTail,
                        int* hostFeedback, int pathTailBelow, cudaStream_t sOut, cudaEvent_t evSplit) {
    const FrameParams& fp = f.fp;
    int launches = 0;
    // Frame pipelining (sOut != s): the end of the frame - the later bounces as k_path_tail, finalize - goes to sOut
    // behind everything on s, so the caller's NEXT frame (other state buffers, enqueued on s) starts under it.
    bool split = false;
    auto out_stream = [&]() -> cudaStream_t {
        if (sOut == nullptr || sOut == s) return s;
        if (!split) {
            cudaEventRecord(evSplit, s);
            cudaStreamWaitEvent(sOut, evSplit, 0);
            split = true;
        }
        return sOut;
    };
    int tailEk = 0;
    size_t tailSmem = 0;
    const bool pathTailOk = path_tail_plan(net, &tailEk, &tailSmem);
    if (fusedTail) tail = nullptr;  // the fused tail kernel needs no graph
    if (tail && !prepare_tail(*tail, f, net, s)) tail = nullptr;
    if (tail) cudaMemcpyAsync(tail->dFrame, &f, sizeof(Frame), cudaMemcpyHostToDevice, s);
    cudaMemsetAsync(f.st.counts, 0, 16 * sizeof(int), s);
    const dim3 tiles((fp.width + 31) / 32, (fp.localHeight + 3) / 4, fp.spp);
    k_entry_camera<<<tiles, kTileThreads, 0, s>>>(f);
    NRB_LAUNCHED("k_entry_camera", s);
    ++launches;
    launches += run_segments(f, net, 0, s, tail, fusedTail);
    k_primary_finish<<<tiles, kTileThreads, 0, s>>>(f);
    NRB_LAUNCHED("k_primary_finish", s);
    ++launches;
    if (lambert) return launches + launch_lambert(f, s);
    const int grid = queue_grid(fp.nElems);
    for (int b = 1; b <= fp.maxBounces; ++b) {
        // Few paths survive the first bounce: the rest of the frame is then one launch.  The decision uses the count
        // the last finished frame reported (hostFeedback, mapped host memory; -1 = not known yet); the kernel is correct
        // for any count, the per-stage kernels below are merely faster for large ones.
        if (b >= 2 && pathTailBelow > 0 && hostFeedback) {
            const int last = *reinterpret_cast<volatile int*>(hostFeedback);
            if ((last < 0 || last <= pathTailBelow) && pathTailOk && launch_path_tail(f, net, b, hostFeedback, out_stream())) {
                ++launches;
                break;
            }
        }
        // path queue 0 -> ray queue 1 (+ segment queue 2) -> path queue 0
        // (k_bounce_sample resets the counters of the bounce: queues 1-3 and the two work cursors)
        k_bounce_sample<<<grid, kQueueThreads, 0, s>>>(f, b, 0, b == 2 ? hostFeedback : nullptr);
        NRB_LAUNCHED("k_bounce_sample", s);
        {
            const int need = (fp.nElems + kQueueThreads - 1) / kQueueThreads;
            const int cap = sm_count() * kBounceCtasPerSm;
            const int g = need < cap ? (need < 1 ? 1 : need) : cap;
            if (f.bounceEntryVote) {
                k_bounce_begin_vote<<<g, kQueueThreads, 0, s>>>(f, b, 0, 1);
            } else {
                const int cap2 = sm_count() * kBounceEntryCtasPerSm;
                k_bounce_begin<<<need < cap2 ? (need < 1 ? 1 : need) : cap2, kQueueThreads, 0, s>>>(f, b, 0, 1);
            }
        }
        NRB_LAUNCHED("k_bounce_begin", s);
        launches += 2;
        bool fusedRan = false;
        launches += run_segments(f, net, 1, s, tail, fusedTail, &fusedRan);
        if (!fusedRan) cudaMemsetAsync(f.st.counts + 0, 0, sizeof(int), s);  // the fused tail kernel resets path queue 0 itself
        k_bounce_finish<<<grid, kQueueThreads, 0, s>>>(f, b, 1, 0);
        NRB_LAUNCHED("k_bounce_finish", s);
        ++launches;
    }
    return launches + launch_finalize(f, out_stream());
}

}  // namespace nrb
