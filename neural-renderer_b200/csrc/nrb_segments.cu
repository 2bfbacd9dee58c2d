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
#include "nrb_frame_dev.cuh"
#include "nrb_neural.cuh"

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
__global__ void __launch_bounds__(kQueueThreads) k_seg_exits(const __grid_constant__ Frame f, int qin) {
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[qin];
    const int D = fp.inputDims;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int e = f.st.queue[qin][i];
        const V3 entry = f4v(f.st.segE[e]), dir = f4v(f.st.segD[e]);
        const V3 shifted = entry + dir * kSegEps;
        const Hit ho = trace_dyn(target_of(f.mesh[NRB_SLOT_OUTER]), TRACE_BACKWARD, shifted, dir);
        const float exitT = ho.t < 1e30f ? ho.t : kSegEps;
        const Hit hi = trace_dyn(target_of(f.mesh[NRB_SLOT_INNER]), TRACE_ANY, shifted, dir);
        const bool hitInner = hi.t < 1e30f;
        const float innerT = hitInner ? hi.t : 1e30f;
        const bool innerFirst = hitInner && (innerT < exitT);
        const V3 exitPos = innerFirst ? shifted + dir * innerT : shifted + dir * exitT;
        f.st.segX[e] = make_float4(exitPos.x, exitPos.y, exitPos.z, exitT);
        f.st.segI[e] = make_float2(innerT, __int_as_float(hitInner ? 1 : 0));
        float* dst = f.st.netIn + static_cast<size_t>(i) * D;
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
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(f.st.stats + 0, 2ull * static_cast<unsigned long long>(n));
        atomicAdd(f.st.stats + 1, static_cast<unsigned long long>(n));
    }
}

// One iteration, second half: decode the network output, record hits, re-enter the shell for the
// rays that go on, and compact them into the next queue.
__global__ void __launch_bounds__(kQueueThreads) k_seg_apply(const __grid_constant__ Frame f, int qin, int qout,
                                                             int recIdx) {
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[qin];
    const int nWarpRounded = (n + 31) & ~31;
    const SurfRec& rec = f.st.rec[recIdx];
    unsigned int nTrace = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nWarpRounded; i += gridDim.x * blockDim.x) {
        bool again = false;
        int e = 0;
        if (i < n) {
            e = f.st.queue[qin][i];
            const uint4 raw = *reinterpret_cast<const uint4*>(f.st.netOut + static_cast<size_t>(i) * 16);
            const float2 o01 = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
            const float2 o23 = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
            const float2 o45 = __half22float2(*reinterpret_cast<const __half2*>(&raw.z));
            const float2 o67 = __half22float2(*reinterpret_cast<const __half2*>(&raw.w));
            const float4 E = f.st.segE[e], X = f.st.segX[e];
            const float2 I = f.st.segI[e];
            const V3 entry = f4v(E), dir = f4v(f.st.segD[e]);
            const float innerT = I.x, exitT = X.w;
            const bool innerHit = __float_as_int(I.y) != 0;
            const bool neuralHit = o01.x >= 0.0f;
            if (neuralHit || (innerHit && innerT < exitT)) {
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
            } else {
                const V3 shiftedExit = f4v(X) + dir * kSegEps;
                const Hit hr = trace_dyn(target_of(f.mesh[NRB_SLOT_OUTER]), TRACE_FORWARD, shiftedExit, dir);
                ++nTrace;
                const bool re = hr.t < 1e30f;
                if (re || innerHit) {
                    const float rd = re ? hr.t : 0.0f;
                    const V3 ne = shiftedExit + dir * rd;
                    f.st.segE[e] = make_float4(ne.x, ne.y, ne.z, E.w + exitT + rd + 2.0f * kSegEps);
                    again = true;
                }
            }
        }
        queue_push(f.st.queue[qout], f.st.counts + qout, again, e);
    }
    stat_flush(f.st.stats + 0, nTrace);
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
__global__ void __launch_bounds__(kQueueThreads) k_bounce_begin(const __grid_constant__ Frame f, int bounce, int qin,
                                                                int qray) {
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[qin];
    const int nWarpRounded = (n + 31) & ~31;
    const SurfRec& cur = f.st.rec[bounce == 1 ? 0 : 1];
    const TraceTarget tOuter = target_of(f.mesh[NRB_SLOT_OUTER]), tInner = target_of(f.mesh[NRB_SLOT_INNER]);
    unsigned int nTrace = 0, nVert = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nWarpRounded; i += gridDim.x * blockDim.x) {
        bool alive = false, enters = false;
        int e = 0, stage = 5;
        V3 org = v3s(0.0f), wi = v3(0.0f, 0.0f, 1.0f), weight = v3s(0.0f);
        float pdf = 0.0f, carried = 0.0f;
        if (i < n) {
            e = f.st.queue[qin][i];
            const ElemId id = queue_elem(fp, e);
            const float4 P = cur.P[e], N = cur.N[e], C = cur.C[e];
            const float2 M = cur.M[e];
            const V3 prev = bounce == 1 ? v3s(0.0f) : f4v(f.st.prevD[e]);
            ++nVert;
            alive = sample_vertex(f, id, f4v(P), f4v(N), f4v(C), v3(C.w, M.x, M.y), bounce == 1, prev, org, wi, weight,
                                  pdf);
            carried = N.w;
        }
        // checkBounceEarlyTerminationKernel + traceOuterShellEntryFromRaysKernel.  The three outer-shell
        // queries of a bounce ray (ANY for early termination, FORWARD entry, BACKWARD exit when the
        // origin is inside the shell) are ONE fused depth-first pass (trace_multi, exact per mode); the
        // inner-shell ANY trace and the re-entry trace then run in a convergent loop over the
        // out-of-line traversal so lanes at different steps still execute the same instructions.
        float tAny = 1e30f, tFwd = 1e30f, tBwd = 1e30f;
        if (alive) {
            trace_multi(tOuter, (carried > 0.0f ? 1 : 0) | 2 | 4, org, wi, tAny, tFwd, tBwd);
            nTrace += (carried > 0.0f ? 1u : 0u) + 1u + (tFwd >= 1e30f ? 1u : 0u);
        }
        const int afterTermination = tFwd < 1e30f ? 5 : (tBwd < 1e30f ? 4 : 5);  // 4 = re-entry trace needed
        if (alive) stage = carried > 0.0f ? 1 : afterTermination;
        Hit hEntry{tFwd, 0.0f, 0.0f, 0u};
        float baseOffset = 0.0f;
        V3 entryOrigin = org;
        while (__any_sync(0xffffffffu, stage < 5)) {
            if (stage < 5) {
                if (stage == 4) {
                    baseOffset = tBwd + kSegEps;
                    entryOrigin = org + wi * baseOffset;
                }
                const TraceTarget tgt = pick_target(stage == 1, tInner, tOuter);
                const Hit h = trace_dyn(tgt, stage == 1 ? TRACE_ANY : TRACE_FORWARD, stage == 4 ? entryOrigin : org, wi);
                ++nTrace;
                if (stage == 1) {
                    if (h.t < 1e30f && tAny < 1e30f && h.t < tAny) { alive = false; stage = 5; }
                    else stage = afterTermination;
                } else {
                    hEntry = h;
                    stage = 5;
                }
            }
        }
        if (alive) {
            f.st.rayO[e] = make_float4(org.x, org.y, org.z, pdf);
            f.st.rayD[e] = make_float4(wi.x, wi.y, wi.z, carried);
            f.st.rayW[e] = make_float4(weight.x, weight.y, weight.z, 0.0f);
            rec_clear(f.st.rec[1], e);
            if (hEntry.t < 1e30f) {
                const V3 p = entryOrigin + wi * hEntry.t;
                f.st.segE[e] = make_float4(p.x, p.y, p.z, baseOffset + hEntry.t);
                f.st.segD[e] = make_float4(wi.x, wi.y, wi.z, 0.0f);
                enters = true;
            }
        }
        queue_push(f.st.queue[qray], f.st.counts + qray, alive, e);
        queue_push(f.st.queue[2], f.st.counts + 2, enters, e);
    }
    stat_flush(f.st.stats + 0, nTrace);
    stat_flush(f.st.stats + 2, nVert);
}

// Bounce, second half: hybrid mesh + merge, throughput/radiance update, Russian roulette.
__global__ void __launch_bounds__(kQueueThreads) k_bounce_finish(const __grid_constant__ Frame f, int bounce, int qray,
                                                                 int qout) {
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[qray];
    const int nWarpRounded = (n + 31) & ~31;
    const SurfRec& rec = f.st.rec[1];
    unsigned int nTrace = 0;
    const bool useAdd = f.additionalEnabled && f.mesh[NRB_SLOT_ADDITIONAL].numTriangles > 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nWarpRounded; i += gridDim.x * blockDim.x) {
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

// The segment loop of traceNeuralSegmentsForRays: fixed kMaxSegIters iterations, device counts.
static int run_segments(const Frame& f, const NetDev& net, int recIdx, cudaStream_t s) {
    int launches = 0;
    int qin = 2;
    for (int it = 0; it < kMaxSegIters; ++it) {
        // Iterations past the first see only rays that left a shell segment without a hit and
        // re-entered it: usually none, rarely many.  They get a one-wave grid (grid-stride loops make
        // any count correct) so that an empty iteration costs a minimal launch.
        const int grid = it == 0 ? queue_grid(f.fp.nElems) : 148;
        const int qout = qin == 2 ? 3 : 2;
        cudaMemsetAsync(f.st.counts + qout, 0, sizeof(int), s);
        k_seg_exits<<<grid, kQueueThreads, 0, s>>>(f, qin);
        NRB_LAUNCHED("k_seg_exits", s);
        launches += 1 + launch_neural_query(net, f.st.netIn, static_cast<size_t>(f.fp.nElems), f.st.counts + qin,
                                            f.st.netOut, s);
        k_seg_apply<<<grid, kQueueThreads, 0, s>>>(f, qin, qout, recIdx);
        NRB_LAUNCHED("k_seg_apply", s);
        ++launches;
        qin = qout;
    }
    return launches;
}

int launch_neural_frame(const Frame& f, const NetDev& net, bool lambert, cudaStream_t s) {
    const FrameParams& fp = f.fp;
    int launches = 0;
    cudaMemsetAsync(f.st.counts, 0, 16 * sizeof(int), s);
    const dim3 tiles((fp.width + 31) / 32, (fp.localHeight + 3) / 4, fp.spp);
    k_entry_camera<<<tiles, kTileThreads, 0, s>>>(f);
    NRB_LAUNCHED("k_entry_camera", s);
    ++launches;
    launches += run_segments(f, net, 0, s);
    k_primary_finish<<<tiles, kTileThreads, 0, s>>>(f);
    NRB_LAUNCHED("k_primary_finish", s);
    ++launches;
    if (lambert) return launches + launch_lambert(f, s);
    const int grid = queue_grid(fp.nElems);
    for (int b = 1; b <= fp.maxBounces; ++b) {
        // path queue 0 -> ray queue 1 (+ segment queue 2) -> path queue 0
        cudaMemsetAsync(f.st.counts + 1, 0, 3 * sizeof(int), s);
        k_bounce_begin<<<grid, kQueueThreads, 0, s>>>(f, b, 0, 1);
        NRB_LAUNCHED("k_bounce_begin", s);
        ++launches;
        launches += run_segments(f, net, 1, s);
        cudaMemsetAsync(f.st.counts + 0, 0, sizeof(int), s);
        k_bounce_finish<<<grid, kQueueThreads, 0, s>>>(f, b, 1, 0);
        NRB_LAUNCHED("k_bounce_finish", s);
        ++launches;
    }
    return launches + launch_finalize(f, s);
}

}  // namespace nrb
