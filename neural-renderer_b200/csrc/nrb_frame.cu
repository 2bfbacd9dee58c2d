// nrb_frame.cu — frame kernels of the ground-truth (classic mesh) path and the shared shading
// stages.  Compiled with -fmad=false (see nrb_math.cuh).
//
// Reference kernels covered here (src/cuda_renderer_neural.cu):
//   intersectGroundTruthKernel :244-296 + initializePathStateKernel :303-354  -> k_gt_primary
//   sampleBounceDirectionsKernel :357-499 + traceGroundTruthBouncesKernel :556-618
//     + integrateBounceKernel :621-675                                        -> k_gt_bounce
//   finalizePathTracingKernel :678-709                                        -> k_finalize
//   lambertKernel :1547-1607                                                  -> k_lambert
// The reference launches each over all W*H*spp elements and stores every intermediate; here one
// bounce is ONE kernel over the compacted queue of live paths, with the sampled direction, BRDF
// weight and pdf living in registers only.
#include "nrb_frame_dev.cuh"

#include <cstdio>
#include <cstdlib>

#include <cub/device/device_radix_sort.cuh>

namespace nrb {

// ---------------------------------------------------------------------------------------------

// Eight resident CTAs per SM (64 registers): 0.367 ms against 0.401 with ten (48 registers) at 1080p.
constexpr int kGtPrimaryCtasPerSm = 8;
__global__ void __launch_bounds__(kTileThreads, kGtPrimaryCtasPerSm) k_gt_primary(const __grid_constant__ Frame f) {
    const FrameParams& fp = f.fp;
    const int total = tile_count(fp);
    unsigned int nRays = 0;
    // Tiles are handed out most-expensive-first (longest processing time first): the cost of a primary tile - silhouette
    // tiles walk hundreds of leaves, sky tiles none - is a property of the camera, so the clocks measured in one frame
    // order the next ones.  With the natural order the launch ended on a few silhouette warps while the SMs were busy
    // 54 % of its duration on average (ncu, profiles/r2_*).
    for (int slot = next_tile(f.st.counts + 10, total); slot >= 0; slot = next_tile(f.st.counts + 10, total)) {
    const int tile = f.st.tileOrder ? f.st.tileOrder[slot] : slot;
    const long long t0 = clock64();
    const ElemId id = tile_elem_at(fp, tile);
    bool active = false;
    V3 dir = v3(0.0f, 0.0f, 1.0f);
    if (id.valid) {
        uint32_t rng = rng_seed(id.pixel, fp.sampleOffset, id.s);
        dir = primary_direction(fp, id.x, id.gy, rng);
    }
    V3 pos = v3s(0.0f), nrm = v3s(0.0f), alb = v3s(0.0f), mp;
    const bool hit = hit_record_sync<TRACE_FORWARD>(f, f.classicSlot, id.valid, fp.camPos, dir, pos, nrm, alb, mp);
    if (id.valid) {
        ++nRays;
        rec_store(f.st.rec[0], id.e, pos, hit ? 1 : 0, nrm, 0.0f, alb, mp);
        V3 L = v3s(0.0f);
        if (!hit) L = fp.useDirectEnvColor ? fp.directEnvColor : env_clamped(f.env, dir, fp.maxRadiance);
        f.st.T[id.e] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
        f.st.L[id.e] = make_float4(L.x, L.y, L.z, 0.0f);
        active = hit && fp.maxBounces > 0;
    }
    queue_push(f.st.queue[0], f.st.counts + 0, active, id.e);
    if (f.st.tileCost && (threadIdx.x & 31) == 0) {
        const long long dt = clock64() - t0;
        f.st.tileCost[tile] = dt > 0xffffffffll ? 0xffffffffu : static_cast<unsigned int>(dt);
    }
    }
    stat_flush(f.st.stats + 0, nRays);
}

// tile ids sorted by descending cost (CUB radix sort, a library primitive); temp storage is grown on demand
__global__ void k_iota(int* v, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = i;
}
size_t tile_sort_temp_bytes(int n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairsDescending(nullptr, bytes, static_cast<const unsigned int*>(nullptr), static_cast<unsigned int*>(nullptr),
                                              static_cast<const int*>(nullptr), static_cast<int*>(nullptr), n);
    return bytes;
}
int launch_tile_sort(const unsigned int* cost, unsigned int* costSorted, int* ids, int* order, int n, void* temp, size_t tempBytes,
                     cudaStream_t s) {
    k_iota<<<(n + 255) / 256, 256, 0, s>>>(ids, n);
    return cub::DeviceRadixSort::SortPairsDescending(temp, tempBytes, cost, costSorted, ids, order, n, 0, 32, s) == cudaSuccess ? 0 : -1;
}

__global__ void __launch_bounds__(kQueueThreads) k_gt_bounce(const __grid_constant__ Frame f, int bounce, int qin,
                                                             int qout, int* hostFeedback) {
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[qin];
    if (hostFeedback && blockIdx.x == 0 && threadIdx.x == 0) {  // paths entering this bounce: the host's hint for the next frames
        hostFeedback[1] = n;
        __threadfence_system();
    }
    const SurfRec& cur = f.st.rec[bounce == 1 ? 0 : 1];
    const SurfRec& nxt = f.st.rec[1];
    unsigned int nTrace = 0, nVert = 0;
    // 32 consecutive vertices per warp at a time from a cursor (the cost of a vertex's ray varies by two orders of
    // magnitude; with a fixed stride the launch ended on its unluckiest warps: 0.437 -> 0.405 ms at 1080p).  This bounce's
    // cursor was cleared by the previous bounce's launch (or by the frame's memset), the next one's is cleared here.
    int* cursor = f.st.counts + 11 + (bounce & 1);
    if (blockIdx.x == 0 && threadIdx.x == 0) f.st.counts[11 + ((bounce + 1) & 1)] = 0;
    for (;;) {
        int base = 0;
        if ((threadIdx.x & 31) == 0) base = atomicAdd(cursor, 32);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const int i = base + (threadIdx.x & 31);
        bool alive = false, sampled = false;
        int e = 0;
        ElemId id{};
        V3 org = v3s(0.0f), wi = v3(0.0f, 0.0f, 1.0f), weight = v3s(0.0f);
        if (i < n) {
            e = f.st.queue[qin][i];
            id = queue_elem(fp, e);
            const float4 P = cur.P[e], N = cur.N[e], C = cur.C[e];
            const float2 M = cur.M[e];
            float pdf;
            const V3 prev = bounce == 1 ? v3s(0.0f) : f4v(f.st.prevD[e]);
            ++nVert;
            sampled = sample_vertex(f, id, f4v(P), f4v(N), f4v(C), v3(C.w, M.x, M.y), bounce == 1, prev, org, wi, weight, pdf);
        }
        {
            // the trace is entered by the whole warp (trace_sync reconverges it every few steps)
            V3 pos = v3s(0.0f), nrm = v3s(0.0f), alb = v3s(0.0f), mp;
            const bool hit = hit_record_sync<TRACE_ANY>(f, f.classicSlot, sampled, org, wi, pos, nrm, alb, mp);
            if (sampled) {
                ++nTrace;
                V3 T = f4v(f.st.T[e]), L = v3s(0.0f);
                bool touchedL;
                if (!hit) L = f4v(f.st.L[e]);
                alive = integrate_vertex(f, id, bounce, hit, wi, weight, T, L, touchedL);
                if (touchedL) f.st.L[e] = make_float4(L.x, L.y, L.z, 0.0f);
                if (alive) {
                    rec_store(nxt, e, pos, 1, nrm, 0.0f, alb, mp);
                    f.st.T[e] = make_float4(T.x, T.y, T.z, 0.0f);
                    f.st.prevD[e] = make_float4(wi.x, wi.y, wi.z, 0.0f);
                }
            }
        }
        queue_push(f.st.queue[qout], f.st.counts + qout, alive, e);
    }
    stat_flush(f.st.stats + 0, nTrace);
    stat_flush(f.st.stats + 2, nVert);
}

// Every ground-truth bounce after the first in ONE launch (the counterpart of k_path_tail): the few paths that survive
// bounce 1 are spread one per warp while they are fewer than the warps, and each is carried through sample -> trace ->
// integrate, bounce after bounce, with its vertex record in registers.  Same arithmetic as k_gt_bounce, so frames are
// bit-identical; what goes away is a kernel boundary (and its latency floor of ~0.1 ms) per bounce.
__global__ void __launch_bounds__(kQueueThreads) k_gt_tail(const __grid_constant__ Frame f, int firstBounce, int qin,
                                                           int* hostFeedback) {
    const FrameParams& fp = f.fp;
    const int n = f.st.counts[qin];
    if (hostFeedback && blockIdx.x == 0 && threadIdx.x == 0) {
        hostFeedback[1] = n;
        __threadfence_system();
    }
    const int lane = threadIdx.x & 31;
    const int totalWarps = (gridDim.x * blockDim.x) >> 5, gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int per = (n + totalWarps - 1) / totalWarps;
    per = per < 1 ? 1 : (per > 32 ? 32 : per);
    unsigned int nTrace = 0, nVert = 0;
    for (int base = gw * per; base < n; base += totalWarps * per) {
        if (lane >= per || base + lane >= n) continue;
        const int e = f.st.queue[qin][base + lane];
        const ElemId id = queue_elem(fp, e);
        const SurfRec& cur = f.st.rec[firstBounce == 1 ? 0 : 1];
        const float4 P = cur.P[e], N = cur.N[e], C = cur.C[e];
        const float2 M = cur.M[e];
        V3 pos = f4v(P), nrm = f4v(N), alb = f4v(C), mp = v3(C.w, M.x, M.y);
        V3 prev = firstBounce == 1 ? v3s(0.0f) : f4v(f.st.prevD[e]);
        V3 T = f4v(f.st.T[e]);
        for (int b = firstBounce; b <= fp.maxBounces; ++b) {
            V3 org, wi, weight;
            float pdf;
            ++nVert;
            if (!sample_vertex(f, id, pos, nrm, alb, mp, b == 1, prev, org, wi, weight, pdf)) break;
            V3 hp, hn, ha, hm;
            const bool hit = hit_record<TRACE_ANY, true>(f, f.classicSlot, org, wi, hp, hn, ha, hm);
            ++nTrace;
            V3 L = v3s(0.0f);
            bool touchedL;
            if (!hit) L = f4v(f.st.L[e]);
            const bool alive = integrate_vertex(f, id, b, hit, wi, weight, T, L, touchedL);
            if (touchedL) f.st.L[e] = make_float4(L.x, L.y, L.z, 0.0f);
            if (!alive) break;
            pos = hp; nrm = hn; alb = ha; mp = hm;
            prev = wi;
        }
    }
    stat_flush(f.st.stats + 0, nTrace);
    stat_flush(f.st.stats + 2, nVert);
}

__global__ void __launch_bounds__(256) k_finalize(const __grid_constant__ Frame f) {
    const FrameParams& fp = f.fp;
    const int lp = blockIdx.x * blockDim.x + threadIdx.x;
    if (lp >= fp.localPixels) return;
    V3 sum = v3s(0.0f);
    for (int s = 0; s < fp.spp; ++s) sum = sum + f4v(f.st.L[lp + s * fp.localPixels]);
    float* a = f.st.accum + 3 * static_cast<size_t>(lp);
    const V3 acc = v3(a[0], a[1], a[2]) + sum;
    a[0] = acc.x; a[1] = acc.y; a[2] = acc.z;
    const float inv = 1.0f / static_cast<float>(fp.sampleOffset + static_cast<uint32_t>(fp.spp));
    const int ly = lp / fp.width, x = lp - ly * fp.width;
    f.st.pixels[global_row(fp, ly) * fp.width + x] = to_srgb8(acc * inv);
}

// bilateralDenoiseKernel (reference src/denoiser.cuh:47-113; shipped there but never launched): joint
// bilateral filter of the running mean, 11x11 window, weights = spatial Gaussian x pow(normal dot, 64)
// x albedo Gaussian, guides = the sample-0 primary-hit record.  Same terms in the same (dy, dx) order as
// the reference loop.  A 32x8 pixel tile and its 5-pixel apron (colour, normal, albedo: 36 B per pixel)
// are staged in shared memory once, so each of the 121 taps is three LDS.96 instead of three global
// reads; the spatial weights are evaluated once per CTA with the same expf.
constexpr int kDnR = 5, kDnTileW = 32, kDnTileH = 8;
constexpr int kDnW = kDnTileW + 2 * kDnR, kDnH = kDnTileH + 2 * kDnR;

__global__ void __launch_bounds__(kDnTileW * kDnTileH) k_denoise(const __grid_constant__ Frame f) {
    const FrameParams& fp = f.fp;
    __shared__ float4 sCol[kDnH][kDnW], sNrm[kDnH][kDnW], sAlb[kDnH][kDnW];  // one LDS.128 per guide and tap
    __shared__ float sSpatial[2 * kDnR + 1][2 * kDnR + 1];
    const int tid = threadIdx.y * kDnTileW + threadIdx.x;
    const int x0 = blockIdx.x * kDnTileW - kDnR, y0 = blockIdx.y * kDnTileH - kDnR;
    const float invAccum = 1.0f / fmaxf(1.0f, static_cast<float>(fp.sampleOffset + static_cast<uint32_t>(fp.spp)));
    for (int i = tid; i < kDnW * kDnH; i += kDnTileW * kDnTileH) {
        const int ty = i / kDnW, tx = i - ty * kDnW;
        const int x = x0 + tx, y = y0 + ty;
        float4 c = make_float4(0.f, 0.f, 0.f, 0.f), n = c, a = c;
        if (x >= 0 && x < fp.width && y >= 0 && y < fp.height) {
            const int p = y * fp.width + x;
            const float* acc = f.st.accum + 3 * static_cast<size_t>(p);
            c = make_float4(acc[0] * invAccum, acc[1] * invAccum, acc[2] * invAccum, 0.f);
            n = f.st.rec[0].N[p];
            a = f.st.rec[0].C[p];
        }
        sCol[ty][tx] = c;
        sNrm[ty][tx] = n;
        sAlb[ty][tx] = a;
    }
    const float invSigmaSpatial2 = 1.0f / (2.0f * 3.0f * 3.0f), invSigmaAlbedo2 = 1.0f / (2.0f * 0.1f * 0.1f);
    if (tid < (2 * kDnR + 1) * (2 * kDnR + 1)) {
        const int dy = tid / (2 * kDnR + 1) - kDnR, dx = tid % (2 * kDnR + 1) - kDnR;
        sSpatial[dy + kDnR][dx + kDnR] = expf(-static_cast<float>(dx * dx + dy * dy) * invSigmaSpatial2);
    }
    __syncthreads();
    const int cx = blockIdx.x * kDnTileW + threadIdx.x, cy = blockIdx.y * kDnTileH + threadIdx.y;
    if (cx >= fp.width || cy >= fp.height) return;
    const int lx = threadIdx.x + kDnR, ly = threadIdx.y + kDnR;
    const float4 cc = sCol[ly][lx], cn = sNrm[ly][lx], ca = sAlb[ly][lx];
    V3 sum = v3s(0.0f);
    float total = 0.0f;
    const int dy0 = max(-kDnR, -cy), dy1 = min(kDnR, fp.height - 1 - cy);
    const int dx0 = max(-kDnR, -cx), dx1 = min(kDnR, fp.width - 1 - cx);
    for (int dy = dy0; dy <= dy1; ++dy) {
        for (int dx = dx0; dx <= dx1; ++dx) {
            const int ty = ly + dy, tx = lx + dx;
            const float4 nn = sNrm[ty][tx], na = sAlb[ty][tx], nc = sCol[ty][tx];
            const float nd = fmaxf(0.0f, fminf(1.0f, cn.x * nn.x + cn.y * nn.y + cn.z * nn.z));
            // pow(nd, 64) as six squarings: nd is in [0, 1], the result differs from powf by < 1e-5 relative
            float wNormal = nd * nd;
            wNormal *= wNormal; wNormal *= wNormal; wNormal *= wNormal; wNormal *= wNormal; wNormal *= wNormal;
            const float ax = na.x - ca.x, ay = na.y - ca.y, az = na.z - ca.z;
            const float wAlbedo = expf(-(ax * ax + ay * ay + az * az) * invSigmaAlbedo2);
            const float w = sSpatial[dy + kDnR][dx + kDnR] * wNormal * wAlbedo;
            sum = sum + v3(nc.x, nc.y, nc.z) * w;
            total += w;
        }
    }
    const V3 color = total > 0.0f ? sum * (1.0f / total) : v3(cc.x, cc.y, cc.z);
    f.st.pixels[cy * fp.width + cx] = to_srgb8(color);
}

__global__ void __launch_bounds__(256) k_lambert(const __grid_constant__ Frame f) {
    const FrameParams& fp = f.fp;
    const int lp = blockIdx.x * blockDim.x + threadIdx.x;
    if (lp >= fp.localPixels) return;
    const int ly = lp / fp.width, x = lp - ly * fp.width;
    const int gy = global_row(fp, ly), pixel = gy * fp.width + x;
    V3 sum = v3s(0.0f);
    for (int s = 0; s < fp.spp; ++s) {
        const int e = lp + s * fp.localPixels;
        uint32_t rng = rng_seed(pixel, fp.sampleOffset, s);
        const V3 dir = primary_direction(fp, x, gy, rng);
        const float4 P = f.st.rec[0].P[e];
        V3 c;
        if (__float_as_int(P.w)) {
            V3 n = f4v(f.st.rec[0].N[e]);
            const float nl = length(n);
            n = nl > 0.0f ? n / nl : v3(0.0f, 1.0f, 0.0f);
            if (dot(n, dir) > 0.0f) n = n * -1.0f;
            c = f4v(f.st.rec[0].C[e]) * fmaxf(0.0f, dot(n, -dir));
        } else {
            c = env_lookup(f.env, dir);
        }
        sum = sum + c;
    }
    f.st.pixels[pixel] = to_srgb8(sum * (1.0f / static_cast<float>(fp.spp)));
}

// traceMeshWithMode on caller rays (the batch entry point nrb_trace_rays).
__global__ void __launch_bounds__(128) k_trace_rays(const __grid_constant__ MeshDev m, int mode, const float* o,
                                                    const float* d, size_t n, Hit* out, unsigned long long* stats) {
    const size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    TraceStats st{0u, 0u};
    if (mode < 3) {  // the closest-hit walk of the frame kernels, entered by the whole warp (trace_sync)
        const bool valid = i < n;
        const V3 ro = valid ? ld3(o, i) : v3s(0.0f), rd = valid ? ld3(d, i) : v3(0.0f, 0.0f, 1.0f);
        Hit h = trace_sync<true>(target_of(m), mode, valid, ro, rd, 1e30f, &st);
        if (h.t >= 1e30f) h = Hit{1e30f, 0.0f, 0.0f, 0u};
        if (valid) out[i] = h;
    } else if (i < n) {
        Hit h;
        if (mode == 3) {  // parity hook for the fused three-mode pass: t = ANY, u = FORWARD, v = BACKWARD distances
            trace_multi(target_of(m), 7, ld3(o, i), ld3(d, i), h.t, h.u, h.v);
            h.prim = 0u;
        } else if (mode == 4) {  // the same through the resumable engine of the bounce entry (union pruning)
            MultiTrace tr;
            tr.begin(target_of(m), 7, ld3(o, i), ld3(d, i));
            while (!tr.done()) {
                const int k = tr.kind();
                if (k == 0) tr.step_node();
                else if (k == 1) tr.step_tri();
                else tr.step_pop();
            }
            h = Hit{tr.best0, tr.best1, tr.best2, 0u};
        } else if (mode >= 5 && mode < 8) {  // parity hook: the run-time-mode loop of the segment stages (cone culling included)
            h = trace_dyn_inline(target_of(m), mode - 5, ld3(o, i), ld3(d, i));
            if (h.t >= 1e30f) h = Hit{1e30f, 0.0f, 0.0f, 0u};
        } else {  // modes 8..10: the first engine (one loop per facing mode), kept as a cross-check of trace_sync
            h = trace_mode<true>(m, mode - 8, ld3(o, i), ld3(d, i), &st);
            if (h.t >= 1e30f) h = Hit{1e30f, 0.0f, 0.0f, 0u};
        }
        out[i] = h;
    }
    if (stats) {
        stat_flush(stats + 3, st.pairs);
        stat_flush(stats + 4, st.tris);
        stat_flush(stats + 0, i < n ? 1u : 0u);
    }
}

// ---------------------------------------------------------------------------------------------

static thread_local StageProfiler* g_profiler = nullptr;
void set_stage_profiler(StageProfiler* p) { g_profiler = p; }
StageProfiler* get_stage_profiler() { return g_profiler; }

bool debug_sync_enabled() {
    static const bool on = getenv("NRB_DEBUG") != nullptr;
    return on;
}

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        cached[dev] = (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) ? n : 148;
    }
    return cached[dev];
}

void debug_check(const char* what, cudaStream_t s) {
    const bool on = debug_sync_enabled();
    if (g_profiler && g_profiler->marks < StageProfiler::kMaxMarks) {
        StageProfiler& p = *g_profiler;
        p.name[p.marks] = what;
        cudaEventRecord(p.ev[++p.marks], s);
    }
    if (!on) return;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    fprintf(stderr, "[nrb] %s: %s\n", what, cudaGetErrorString(e));
}

void launch_trace_rays(const MeshDev& m, int mode, const float* o, const float* d, size_t n, Hit* out,
                       unsigned long long* stats, cudaStream_t s) {
    if (n == 0) return;
    k_trace_rays<<<static_cast<unsigned>((n + 127) / 128), 128, 0, s>>>(m, mode, o, d, n, out, stats);
    NRB_LAUNCHED("k_trace_rays", s);
}

int launch_finalize(const Frame& f, cudaStream_t s) {
    k_finalize<<<(f.fp.localPixels + 255) / 256, 256, 0, s>>>(f);
    NRB_LAUNCHED("k_finalize", s);
    return 1;
}

// ---- frame gather over peer memory (multi-GPU) -----------------------------------------------------
// With nrb_gather_attach a non-root rank's k_finalize stores its row bands straight into the root GPU's
// frame buffer (NVLink peer stores); these three single-thread kernels are the hand-shake around it.
// sync[0] = bands-complete signals received (monotonic), sync[1] = frames the root has released for
// overwriting, both in the root's memory; err = local flag set when a wait gives up.
__device__ __forceinline__ unsigned ld_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__global__ void k_gather_release(unsigned* sync, unsigned frames) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(sync + 1), "r"(frames) : "memory");
}
__global__ void k_gather_wait(const unsigned* flag, unsigned want, unsigned* err) {
    const long long t0 = clock64();
    while (static_cast<int>(ld_sys(flag) - want) < 0) {
        if (clock64() - t0 > 6000000000ll) {  // ~3 s: a peer stopped rendering; give up instead of hanging the GPU
            *err = 1u;
            break;
        }
        __nanosleep(500);
    }
    __threadfence_system();
}
__global__ void k_gather_signal(unsigned* remoteSync) {
    __threadfence_system();  // the finalize kernel before this one wrote the bands; order them before the signal
    atomicAdd_system(remoteSync, 1u);
}
void launch_gather_release(unsigned* sync, unsigned frames, cudaStream_t s) {
    k_gather_release<<<1, 1, 0, s>>>(sync, frames);
    NRB_LAUNCHED("k_gather_release", s);
}
void launch_gather_wait(const unsigned* flag, unsigned want, unsigned* err, cudaStream_t s) {
    k_gather_wait<<<1, 1, 0, s>>>(flag, want, err);
    NRB_LAUNCHED("k_gather_wait", s);
}
void launch_gather_signal(unsigned* remoteSync, cudaStream_t s) {
    k_gather_signal<<<1, 1, 0, s>>>(remoteSync);
    NRB_LAUNCHED("k_gather_signal", s);
}

int launch_denoise(const Frame& f, cudaStream_t s) {
    const dim3 grid((f.fp.width + kDnTileW - 1) / kDnTileW, (f.fp.height + kDnTileH - 1) / kDnTileH);
    k_denoise<<<grid, dim3(kDnTileW, kDnTileH), 0, s>>>(f);
    NRB_LAUNCHED("k_denoise", s);
    return 1;
}

int launch_lambert(const Frame& f, cudaStream_t s) {
    k_lambert<<<(f.fp.localPixels + 255) / 256, 256, 0, s>>>(f);
    NRB_LAUNCHED("k_lambert", s);
    return 1;
}

int launch_gt_frame(const Frame& f, bool lambert, cudaStream_t s, int* hostFeedback, int pathTailBelow, cudaStream_t sOut,
                    cudaEvent_t evSplit) {
    const FrameParams& fp = f.fp;
    int launches = 0;
    // frame pipelining (sOut != s): the end of the frame (k_gt_tail, finalize) goes to sOut behind everything on s
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
    cudaMemsetAsync(f.st.counts, 0, 16 * sizeof(int), s);
    {
        const int tiles = ((fp.width + 7) / 8) * ((fp.localHeight + 3) / 4) * fp.spp;
        const int need = (tiles + 3) / 4, cap = sm_count() * kGtPrimaryCtasPerSm;
        k_gt_primary<<<need < cap ? need : cap, kTileThreads, 0, s>>>(f);
    }
    NRB_LAUNCHED("k_gt_primary", s);
    ++launches;
    if (lambert) return launches + launch_lambert(f, s);
    int qin = 0;
    for (int b = 1; b <= fp.maxBounces; ++b) {
        const int qout = qin ^ 1;
        if (b >= 2 && pathTailBelow > 0 && hostFeedback) {  // few paths left: the rest of the frame is one launch (see k_path_tail)
            const int last = reinterpret_cast<volatile int*>(hostFeedback)[1];
            if (last < 0 || last <= pathTailBelow) {
                cudaStream_t so = out_stream();
                k_gt_tail<<<sm_count() * 8, kQueueThreads, 0, so>>>(f, b, qin, hostFeedback);
                NRB_LAUNCHED("k_gt_tail", so);
                ++launches;
                break;
            }
        }
        cudaMemsetAsync(f.st.counts + qout, 0, sizeof(int), s);
        k_gt_bounce<<<queue_grid(fp.nElems), kQueueThreads, 0, s>>>(f, b, qin, qout, b == 2 ? hostFeedback : nullptr);
        NRB_LAUNCHED("k_gt_bounce", s);
        ++launches;
        qin = qout;
    }
    return launches + launch_finalize(f, out_stream());
}

}  // namespace nrb
