// nrb_api.cu — the C ABI of libnrb200.so (include/nrb200.h) and the host orchestration behind it.
//
// Mirrors the host side of the reference's RendererNeural (src/cuda_renderer_neural.cu:1632-3003):
// constructor/network shape, weight loading, buffer lifecycle, accumulation-reset rules and the
// render() entry point.  The frame itself is enqueued by nrb_frame.cu / nrb_segments.cu /
// nrb_neural.cu on one non-blocking stream per renderer and joined to the legacy default stream
// before nrb_render returns (callers issue a plain cudaMemcpy from devicePixels() next,
// src/evaluate.cu:266).  There is no CPU fallback anywhere in this file: every entry point that
// needs a device fails with NRB_ERR_CUDA when none is usable.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/nrb200.h"
#include "nrb_frame.cuh"
#include "nrb_neural.cuh"
#include "nrb_trace.cuh"

using namespace nrb;

static_assert(sizeof(nrb_material) == sizeof(MaterialDev), "material layout");
static_assert(sizeof(nrb_hit) == sizeof(Hit), "hit layout");
static_assert(sizeof(nrb_bvh_node) == 44, "BvhNode layout");

namespace {

thread_local std::string g_error;

int fail(int code, const std::string& msg) {
    g_error = msg;
    return code;
}

#define NRB_CUDA(expr)                                                                              \
    do {                                                                                            \
        cudaError_t err__ = (expr);                                                                 \
        if (err__ != cudaSuccess)                                                                   \
            return fail(NRB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(err__));       \
    } while (0)

// Owning device allocation.
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    bool borrowed = false;  // an alias of another renderer's allocation (frame slices): never freed here
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), bytes(o.bytes), borrowed(o.borrowed) { o.p = nullptr; o.bytes = 0; o.borrowed = false; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        if (this != &o) { release(); p = o.p; bytes = o.bytes; borrowed = o.borrowed; o.p = nullptr; o.bytes = 0; o.borrowed = false; }
        return *this;
    }
    ~DevBuf() { release(); }
    void release() {
        if (p && !borrowed) cudaFree(p);
        p = nullptr;
        bytes = 0;
        borrowed = false;
    }
    void alias(void* ptr, size_t n) {
        release();
        p = ptr;
        bytes = n;
        borrowed = ptr != nullptr;
    }
    cudaError_t alloc(size_t n) {
        release();
        if (n == 0) return cudaSuccess;
        cudaError_t e = cudaMalloc(&p, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    cudaError_t upload(const void* src, size_t n, cudaStream_t s) {
        cudaError_t e = alloc(n);
        if (e != cudaSuccess || n == 0) return e;
        e = cudaMemcpyAsync(p, src, n, cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) return e;
        return cudaStreamSynchronize(s);  // the host array is borrowed for the call only
    }
    template <class T>
    T* as() const { return static_cast<T*>(p); }
};

struct MeshSlot {
    DevBuf nodes, tris, vertices, normals, texcoords, indices, materials, materialMap, materialIds, textures;
    std::vector<DevBuf> texPixels;
    MeshDev view{};
    float boundsMin[3] = {0, 0, 0}, boundsMax[3] = {0, 0, 0};
    int depth = 0;
    bool closed = false;  // closed, consistently oriented surface (computed for the outer shell only)
    void clear() {
        nodes.release(); tris.release(); vertices.release(); normals.release(); texcoords.release();
        indices.release(); materials.release(); materialMap.release(); materialIds.release(); textures.release();
        texPixels.clear();
        view = MeshDev{};
        depth = 0;
        closed = false;
    }
};

}  // namespace

struct nrb_renderer {
    int device = 0;
    // Frame slices (NRB_OPT_FRAME_SLICES = S > 1): the rows this renderer owns are dealt to S slice renderers - children
    // with their own wavefront state, queues and stream, borrowing this renderer's scene, weights and frame buffer -
    // through the very tile mechanism of the multi-GPU partition (virtual rank = rank + k * world of world * S).  Their
    // stage chains run concurrently, so the latency-bound tails of one slice (the long rays that end a traversal
    // launch, the later bounces) run under the throughput stages of another.  Pixels are those of the unsliced frame.
    int slices = 0;        // NRB_OPT_FRAME_SLICES: 0 = automatic (4 in ground-truth mode, 1 in neural mode), else the count
    int activeSlices = 1;  // what the current frame uses
    std::vector<nrb_renderer*> kids;
    nrb_renderer* parent = nullptr;
    cudaEvent_t evFork = nullptr, evDone = nullptr;
    uchar4* pixelsOverride = nullptr;
    bool lastSliced = false;
    cudaStream_t stream = nullptr;
    cudaEvent_t evStart = nullptr, evStop = nullptr, evJoin = nullptr, evJoinIn = nullptr;
    // Frame pipelining (NRB_OPT_PIPELINE, neural mode): the frame's stages up to the end of bounce 1 run on `early`, its
    // end (k_path_tail, finalize) on `stream` behind them, and consecutive frames alternate between two wavefront-state
    // buffers - so frame k+1 starts under the latency-bound end of frame k.  `stream` stays the stream every frame
    // COMPLETES on, in order (read-backs, gather, nrb_sync and callers' events on nrb_stream() keep their meaning); work
    // other entry points enqueue on it between frames is waited for by the next frame's early part (`touched`).
    int pipeline = 1;
    cudaStream_t early = nullptr;
    cudaEvent_t evSplit = nullptr, evTouch = nullptr, evFrameDone[2] = {nullptr, nullptr};
    bool frameDoneValid[2] = {false, false};
    bool touched = true;   // something other than a pipelined frame was enqueued on `stream` since the last frame
    int parity = 0;        // which state buffer the last frame used
    State stBuf[2];
    bool stBufValid[2] = {false, false};
    // nrb_read_pixels_async: frame -> staging copy on `stream`, staging -> host on `copyStream`
    cudaStream_t copyStream = nullptr;
    cudaEvent_t evStaged = nullptr, evCopied = nullptr;
    bool copyPending = false;
    // peer-memory frame gather (nrb_gather_export / nrb_gather_attach)
    struct Gather {
        int role = 0;                 // 0 off, 1 root (owns the gathered frame), 2 peer (stores into the root's)
        int peers = 0;                // root: ranks that signal per frame
        unsigned frames = 0;          // frames rendered since the gather was set up
        DevBuf sync;                  // root: {signals, released, -, -}; every rank: [4] = local error flag
        void* remotePixels = nullptr; // peer: the root's frame buffer and sync block, opened over CUDA IPC
        void* remoteSync = nullptr;
    } gather;

    // scene
    MeshSlot mesh[4];
    DevBuf envPixels;
    int envW = 0, envH = 0;
    float envStrength = 1.0f;
    MaterialDev globalMaterial{};

    // network
    nrb_nn_config cfg{};
    NetDev net{};
    DevBuf grid, mlp, packedW, xt, encCur;
    SegTail segTail;
    int tailMode = 2;  // NRB_OPT_TAIL_GRAPH: 0 fixed loop, 1 conditional WHILE graph, 2 fused tail kernel
    int pathTailBelow = 49152;  // NRB_OPT_PATH_TAIL: bounces >= 2 run as one kernel while they start with at most this many paths
    int* pathFeedback = nullptr;  // mapped host memory: [0] paths that entered bounce 2 in the last finished frame
    bool denoise = false;
    bool literalBounceTraces = false;
    int bounceEntry = 0;  // NRB_OPT_BOUNCE_ENTRY
    size_t gridParams = 0, mlpParams = 0;
    bool weightsLoaded = false;

    // RendererNeural state (src/cuda_renderer_neural.h:216-277)
    float fwd[3] = {0, 0, -1}, right[3] = {1, 0, 0}, up[3] = {0, 1, 0}, fovY = 0.785f;
    int width = 0, height = 0;
    bool lambertView = false, useNeuralQuery = false;
    int samplesPerPixel = 1, bounceCount = 0, classicMeshIndex = 0;
    float envmapRotation = 0.0f;
    bool useConstantNeuralColor = false, useDirectEnvColor = false, useAdditionalMesh = true;
    float constantNeuralColor[3] = {0.8f, 0.8f, 0.8f}, directEnvColor[3] = {0, 0, 0};
    uint32_t accumSampleCount = 0;
    bool lastUseNeuralQuery = true, lastLambertView = false, lastUseAdditionalMesh = true, hasLastCamera = false;
    int lastBounceCount = -1, lastSamplesPerPixel = -1, lastClassicMeshIndex = -1;
    float lastEnvmapRotation = 0.0f, lastCam[3] = {0, 0, 0}, lastFwd[3], lastRight[3], lastUp[3], lastFovY = 0.0f;
    int capture = 0;
    int bandRows = 16, rank = 0, world = 1;

    // wavefront state
    std::vector<DevBuf> pool;
    State st{};
    DevBuf accum, pixels, pixelsStaged;
    // primary-pass tile order (ground-truth mode): costs of a frame, ids sorted by them, sort scratch
    DevBuf tileCost, tileCostSorted, tileIds, tileOrder, tileSortTemp;
    int tileCount = 0;
    int tileOrderState = 0;  // 0 none, 2 order valid, 3 order usable but to be re-sorted from this frame's costs (camera moved)
    bool tileOrderOn = true; // NRB_OPT_TILE_ORDER
    size_t stateElems = 0, accumPixels = 0, pixelCount = 0;
    int localHeight = 0;

    // last frame
    double lastMs = 0.0;
    bool lastNeural = false;
    int lastRayCount = 0, lastLaunches = 0;
    uint64_t lastStats[8] = {0};
    DevBuf scratchA, scratchB, scratchC;  // host-variant batch entry points

    // per-kernel device time (NRB_OPT_PROFILE)
    int profile = 0;
    StageProfiler prof;
    bool profPending = false;
    std::map<std::string, std::pair<double, uint64_t>> stages;  // name -> (total ms, launches)
    std::vector<std::string> stageOrder;
    double lastBuckets[NRB_TIMING_BUCKETS] = {0};  // the last profiled frame in the reference's KernelId buckets
    bool lastBucketsValid = false;
};

namespace {

void gather_reset(nrb_renderer* r) {
    if (r->gather.remotePixels) cudaIpcCloseMemHandle(r->gather.remotePixels);
    if (r->gather.remoteSync) cudaIpcCloseMemHandle(r->gather.remoteSync);
    r->gather.remotePixels = r->gather.remoteSync = nullptr;
    r->gather.role = 0;
    r->gather.peers = 0;
    r->gather.frames = 0;
}

int local_rows(const nrb_renderer* r) {
    if (r->world <= 1) return r->height;
    int rows = 0;
    for (int y0 = r->rank * r->bandRows; y0 < r->height; y0 += r->world * r->bandRows)
        rows += std::min(r->bandRows, r->height - y0);
    return rows;
}

// ---- network shape (RendererNeural ctor :1647-1688; GridEncodingTemplated ctor, tiny-cuda-nn
// encodings/grid.h:673-737; grid_scale / grid_resolution, common_device.h:886-895) --------------
int configure_network(nrb_renderer* r, const nrb_nn_config* c) {
    nrb_nn_config d{};
    if (c) d = *c;
    if (d.log2_hashmap_size <= 0) d.log2_hashmap_size = 14;
    if (d.base_resolution <= 0) d.base_resolution = 16;
    if (d.n_points <= 0) d.n_points = 3;
    if (d.n_levels <= 0) d.n_levels = 8;
    if (d.n_features_per_level <= 0) d.n_features_per_level = 4;
    if (d.per_level_scale <= 0.0f) d.per_level_scale = 2.0f;
    if (d.sh_degree == 0) d.sh_degree = 4;
    else if (d.sh_degree < 0) d.sh_degree = 0;
    if (d.mlp_width <= 0) d.mlp_width = 128;
    if (d.n_hidden_layers <= 0) d.n_hidden_layers = 4;
    if (d.n_levels > kMaxLevels || d.n_points > 4 || d.sh_degree > 4 || d.log2_hashmap_size > 28)
        return fail(NRB_ERR_INVALID, "network shape out of range");
    r->cfg = d;
    NetDev& n = r->net;
    n = NetDev{};
    n.nPoints = d.n_points; n.nLevels = d.n_levels; n.nFeatures = d.n_features_per_level;
    n.shDegree = d.sh_degree; n.width = d.mlp_width; n.nHidden = d.n_hidden_layers;
    n.inputDims = 3 * d.n_points + (d.sh_degree > 0 ? 3 : 0);
    n.encDims = d.n_points * d.n_levels * d.n_features_per_level + d.sh_degree * d.sh_degree;
    const float log2Scale = std::log2(d.per_level_scale);
    uint32_t offset = 0;
    for (int l = 0; l < d.n_levels; ++l) {
        const float scale = exp2f(l * log2Scale) * d.base_resolution - 1.0f;
        const uint32_t res = static_cast<uint32_t>(ceilf(scale)) + 1u;
        const uint32_t cap = 0xFFFFFFFFu / 2u;
        uint32_t entries = std::pow(static_cast<float>(res), 3) > static_cast<float>(cap) ? cap : res * res * res;
        entries = (entries + 7u) / 8u * 8u;
        entries = std::min(entries, 1u << d.log2_hashmap_size);
        n.levelOffset[l] = offset;
        n.levelSize[l] = entries;
        n.levelRes[l] = res;
        n.levelScale[l] = scale;
        // grid_index: linear while res^3 fits the level, else hashed (stride overflow guard at 0x659)
        const bool fits = res <= 0x659u && static_cast<uint64_t>(res) * res * res <= entries;
        n.levelDense[l] = fits ? 1 : 0;
        n.levelPow2[l] = (entries & (entries - 1u)) == 0 ? 1 : 0;
        offset += entries;
    }
    r->gridParams = static_cast<size_t>(offset) * d.n_features_per_level;
    const size_t w = d.mlp_width;
    r->mlpParams = w * n.encDims + static_cast<size_t>(d.n_hidden_layers - 1) * w * w + 16 * w;
    int warps = 0;
    if (neural_query_smem(n, &warps) == 0) return fail(NRB_ERR_INVALID, "unsupported network shape for the fused query kernel");
    return NRB_OK;
}

// Is the mesh a closed, consistently oriented surface?  Vertices are welded by exact position; every directed edge
// must occur once and its reverse once.  A ray that starts inside such a shell always has a back face ahead of it,
// which is what lets k_seg_exits stop the outer-shell exit trace at the inner-shell hit; for any other
// outer shell the exit trace runs to the end, exactly as the reference's does.
bool closed_surface(const float* vertices, int nv, const uint32_t* indices, int nt) {
    if (nt < 4) return false;
    std::vector<uint32_t> order(nv), canon(nv);
    for (int i = 0; i < nv; ++i) order[i] = static_cast<uint32_t>(i);
    auto key = [&](uint32_t i) { return vertices + 3 * static_cast<size_t>(i); };
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return std::memcmp(key(a), key(b), 12) < 0; });
    for (int i = 0; i < nv; ++i)
        canon[order[i]] = (i > 0 && std::memcmp(key(order[i]), key(order[i - 1]), 12) == 0) ? canon[order[i - 1]] : order[i];
    std::vector<uint64_t> edges;
    edges.reserve(3 * static_cast<size_t>(nt));
    for (int t = 0; t < nt; ++t) {
        const uint32_t v[3] = {canon[indices[3 * t]], canon[indices[3 * t + 1]], canon[indices[3 * t + 2]]};
        if (v[0] == v[1] || v[1] == v[2] || v[2] == v[0]) return false;
        for (int k = 0; k < 3; ++k) edges.push_back((static_cast<uint64_t>(v[k]) << 32) | v[(k + 1) % 3]);
    }
    std::sort(edges.begin(), edges.end());
    for (size_t i = 1; i < edges.size(); ++i)
        if (edges[i] == edges[i - 1]) return false;
    for (uint64_t e : edges)
        if (!std::binary_search(edges.begin(), edges.end(), (e << 32) | (e >> 32))) return false;
    return true;
}

// ---- BVH re-layout ------------------------------------------------------------------------------
// Normal cones of every node's subtree (nrb_scene.cuh: cone word).  Axis = the quantised mean normal of the subtree's
// triangles, threshold from the EXACT minimum of axis . n over those triangles (double precision, the fp32 normals the
// kernels' facing test uses), widened by the margins described below.  O(triangles x depth); shells only.
std::vector<uint32_t> node_cones(const nrb_bvh_node* src, int nn, const std::vector<TriBlock>& tris) {
    std::vector<uint32_t> cone(nn, kNoCone);
    struct D3 { double x, y, z; };
    std::vector<D3> sum(nn, D3{0, 0, 0});
    auto normal = [&](int t, D3& n) {
        n = D3{tris[t].p.w, tris[t].q.w, tris[t].r.w};
        const double l = std::sqrt(n.x * n.x + n.y * n.y + n.z * n.z);
        if (!(l > 0.5) || !std::isfinite(l)) return false;  // degenerate triangle: its normal is 0 or NaN, it can never be hit
        n = D3{n.x / l, n.y / l, n.z / l};
        return true;
    };
    // post-order over the tree (children have larger indices than their parent in every builder we accept? not
    // guaranteed: use an explicit order)
    std::vector<int> order;
    order.reserve(nn);
    {
        std::vector<int> todo{0};
        while (!todo.empty()) {
            const int i = todo.back();
            todo.pop_back();
            order.push_back(i);
            if (!src[i].is_leaf) { todo.push_back(src[i].left); todo.push_back(src[i].right); }
        }
    }
    for (auto it = order.rbegin(); it != order.rend(); ++it) {
        const int i = *it;
        if (src[i].is_leaf) {
            for (int t = src[i].first; t < src[i].first + src[i].count; ++t) {
                D3 n;
                if (normal(t, n)) { sum[i].x += n.x; sum[i].y += n.y; sum[i].z += n.z; }
            }
        } else {
            const D3 &a = sum[src[i].left], &b = sum[src[i].right];
            sum[i] = D3{a.x + b.x, a.y + b.y, a.z + b.z};
        }
    }
    std::vector<int> stack;
    for (int i = 0; i < nn; ++i) {
        const D3 s = sum[i];
        const double l = std::sqrt(s.x * s.x + s.y * s.y + s.z * s.z);
        if (!(l > 1e-9)) continue;
        const int qx = static_cast<int>(std::lround(s.x / l * 127.0)), qy = static_cast<int>(std::lround(s.y / l * 127.0)),
                  qz = static_cast<int>(std::lround(s.z / l * 127.0));
        const double ax = qx / 127.0, ay = qy / 127.0, az = qz / 127.0, al = std::sqrt(ax * ax + ay * ay + az * az);
        if (!(al > 0.5)) continue;
        double minDot = 1.0;
        stack.assign(1, i);
        while (!stack.empty()) {
            const int k = stack.back();
            stack.pop_back();
            if (src[k].is_leaf) {
                for (int t = src[k].first; t < src[k].first + src[k].count; ++t) {
                    D3 n;
                    if (normal(t, n)) minDot = std::min(minDot, (ax * n.x + ay * n.y + az * n.z) / al);
                }
            } else {
                stack.push_back(src[k].left);
                stack.push_back(src[k].right);
            }
        }
        // every normal is within theta of the unit axis; n . d >= eps for all of them once axis . d >= sin(theta + asin eps).
        // eps = 1e-5 keeps the sign of the kernels' fp32 facing dot (rounding ~ 4e-7); the kernel tests the UNnormalised
        // quantised axis against s * |d|, so s carries |axis|, the fp32 error of that dot and of |d| (4e-6), is rounded
        // up to the next 1/255 and given one more step.
        const double theta = std::acos(std::max(-1.0, std::min(1.0, minDot))) + std::asin(1e-5);
        if (!(theta < 1.5707)) continue;
        const int sq = static_cast<int>(std::ceil((std::sin(theta) * al + 4e-6) * 255.0)) + 1;
        if (sq >= 255) continue;
        cone[i] = (static_cast<uint32_t>(qx) & 0xffu) | ((static_cast<uint32_t>(qy) & 0xffu) << 8) |
                  ((static_cast<uint32_t>(qz) & 0xffu) << 16) | (static_cast<uint32_t>(sq) << 24);
    }
    return cone;
}

int relayout_mesh(nrb_renderer* r, MeshSlot& m, const nrb_mesh_desc* d, bool withCones) {
    const int nn = d->num_nodes, nt = d->num_triangles;
    const nrb_bvh_node* src = d->nodes;
    if (!src || nn <= 0 || !d->vertices || !d->indices) return fail(NRB_ERR_INVALID, "mesh without nodes/vertices/indices");
    for (int i = 0; i < 3 * nt; ++i)
        if (d->indices[i] >= static_cast<uint32_t>(d->num_vertices)) return fail(NRB_ERR_INVALID, "vertex index out of range");

    // triangle blocks in leaf (= primitive id) order; same operation order as the reference's
    // per-candidate normalize(cross(v1 - v0, v2 - v0)) (src/mesh_traversal.cuh:163)
    std::vector<TriBlock> tris(nt);
    for (int i = 0; i < nt; ++i) {
        const V3 v0 = ld3(d->vertices, d->indices[3 * i]), v1 = ld3(d->vertices, d->indices[3 * i + 1]),
                 v2 = ld3(d->vertices, d->indices[3 * i + 2]);
        const V3 e1 = v1 - v0, e2 = v2 - v0;
        const V3 n = normalize(cross(e1, e2));
        tris[i].p = make_float4(v0.x, v0.y, v0.z, n.x);
        tris[i].q = make_float4(e1.x, e1.y, e1.z, n.y);
        tris[i].r = make_float4(e2.x, e2.y, e2.z, n.z);
    }

    auto check_leaf = [&](const nrb_bvh_node& n) {
        return n.count >= 1 && n.first >= 0 && n.first + n.count <= nt && n.first + n.count <= (1 << (31 - kLeafCountBits));
    };
    std::vector<NodePair> pairs;
    MeshDev& v = m.view;
    int extraDepth = 0;
    // node indices are validated below before they are followed; the cone pass needs a well-formed tree, so it runs
    // only after a first structural check
    std::vector<uint32_t> cones;
    if (withCones) {
        bool ok = true;
        std::vector<char> seenNode(nn, 0);
        std::vector<int> todo{0};
        seenNode[0] = 1;
        while (ok && !todo.empty()) {
            const int i = todo.back();
            todo.pop_back();
            if (src[i].is_leaf) { ok = check_leaf(src[i]); continue; }
            for (int c : {src[i].left, src[i].right}) {
                if (c <= 0 || c >= nn || seenNode[c]) { ok = false; break; }
                seenNode[c] = 1;
                todo.push_back(c);
            }
        }
        if (ok) cones = node_cones(src, nn, tris);
    }
    auto cone_of = [&](int node) { return cones.empty() ? kNoCone : cones[node]; };
    // intersectAabb pads every box by 1e-10 before the slab test (src/mesh_traversal.cuh:67-69); the
    // padded boxes are stored so the kernels skip 12 additions per node pair (same fp32 operations)
    const float kPad = 1e-10f;
    auto lo = [&](float x) { return x - kPad; };
    auto hi = [&](float x) { return x + kPad; };
    for (int k = 0; k < 3; ++k) {
        v.rootMin[k] = lo(src[0].bounds_min[k]);
        v.rootMax[k] = hi(src[0].bounds_max[k]);
    }
    // A leaf reference holds at most 32 triangles.  The reference accepts any leaf size (src/mesh_traversal.cuh:155-185):
    // a larger leaf becomes a chain of node pairs whose children all carry the LEAF'S box - the first 32 triangles on
    // the left, the rest on the right.  Equal entry distances visit left first, so the triangles are still tested in
    // ascending order; the only difference is that the box is compared with the current best hit again between blocks
    // (a block is skipped only when the box lies entirely beyond a hit already found, which no triangle inside it
    // could beat).
    auto emit_leaf = [&](const nrb_bvh_node& n) -> int {
        const int kMax = 1 << kLeafCountBits;
        if (n.count <= kMax) return leaf_ref(n.first, n.count);
        const float4 xy = make_float4(lo(n.bounds_min[0]), lo(n.bounds_min[1]), hi(n.bounds_max[0]), hi(n.bounds_max[1]));
        const float2 z = make_float2(lo(n.bounds_min[2]), hi(n.bounds_max[2]));
        const int head = static_cast<int>(pairs.size());
        int chain = 0;
        for (int first = n.first, left = n.count; left > kMax; first += kMax, left -= kMax, ++chain) {
            const int rest = left - kMax;
            NodePair p;
            p.a = xy; p.b = xy; p.c = make_float4(z.x, z.y, z.x, z.y);
            const int self = static_cast<int>(pairs.size());
            const int rightRef = rest <= kMax ? leaf_ref(first + kMax, rest) : self + 1;
            p.d = make_int4(leaf_ref(first, kMax), rightRef, static_cast<int>(kNoCone), static_cast<int>(kNoCone));
            pairs.push_back(p);
        }
        extraDepth = std::max(extraDepth, chain);
        return head;
    };
    int maxDepth = 1;
    if (src[0].is_leaf) {
        if (!check_leaf(src[0])) return fail(NRB_ERR_INVALID, "bad root leaf");
        v.rootRef = emit_leaf(src[0]);
    } else {
        // DFS pre-order numbering of inner nodes keeps a subtree's pairs adjacent in memory
        struct Item { int node, pair, depth; };
        std::vector<Item> todo;
        pairs.reserve(nn / 2 + 1);
        pairs.emplace_back();
        todo.push_back({0, 0, 1});
        std::vector<char> seen(nn, 0);
        while (!todo.empty()) {
            const Item it = todo.back();
            todo.pop_back();
            const nrb_bvh_node& n = src[it.node];
            maxDepth = std::max(maxDepth, it.depth);
            const int kids[2] = {n.left, n.right};
            int refs[2];
            for (int k = 0; k < 2; ++k) {
                const int c = kids[k];
                if (c <= 0 || c >= nn || seen[c]) return fail(NRB_ERR_INVALID, "bad child index in BVH");
                seen[c] = 1;
                if (src[c].is_leaf) {
                    if (!check_leaf(src[c])) return fail(NRB_ERR_INVALID, "leaf range out of bounds");
                    refs[k] = emit_leaf(src[c]);
                } else {
                    refs[k] = static_cast<int>(pairs.size());
                    pairs.emplace_back();
                }
            }
            // right first on the stack so the left subtree is numbered first
            if (!src[kids[1]].is_leaf) todo.push_back({kids[1], refs[1], it.depth + 1});
            if (!src[kids[0]].is_leaf) todo.push_back({kids[0], refs[0], it.depth + 1});
            const nrb_bvh_node& L = src[kids[0]];
            const nrb_bvh_node& R = src[kids[1]];
            NodePair& p = pairs[it.pair];
            p.a = make_float4(lo(L.bounds_min[0]), lo(L.bounds_min[1]), hi(L.bounds_max[0]), hi(L.bounds_max[1]));
            p.b = make_float4(lo(R.bounds_min[0]), lo(R.bounds_min[1]), hi(R.bounds_max[0]), hi(R.bounds_max[1]));
            p.c = make_float4(lo(L.bounds_min[2]), hi(L.bounds_max[2]), lo(R.bounds_min[2]), hi(R.bounds_max[2]));
            p.d = make_int4(refs[0], refs[1], static_cast<int>(cone_of(kids[0])), static_cast<int>(cone_of(kids[1])));
        }
        v.rootRef = 0;
    }
    maxDepth += extraDepth;
    if (maxDepth >= kTraceStack) return fail(NRB_ERR_INVALID, "BVH deeper than the traversal stack (64)");
    m.depth = maxDepth;
    NRB_CUDA(m.nodes.upload(pairs.data(), pairs.size() * sizeof(NodePair), r->stream));
    NRB_CUDA(m.tris.upload(tris.data(), tris.size() * sizeof(TriBlock), r->stream));
    v.nodes = m.nodes.as<NodePair>();
    v.tris = m.tris.as<TriBlock>();
    v.numPairs = static_cast<int>(pairs.size());
    v.numTriangles = nt;
    return NRB_OK;
}

}  // namespace

// =================================================================================================
// C entry points
// =================================================================================================
extern "C" {

const char* nrb_last_error(void) { return g_error.c_str(); }
int nrb_abi_version(void) { return NRB_ABI_VERSION; }

int nrb_create(const nrb_nn_config* cfg, int device, nrb_renderer** out) {
    if (!out) return fail(NRB_ERR_INVALID, "out is null");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(NRB_ERR_CUDA, std::string("no usable CUDA device (libnrb200 has no CPU path): ") +
                                      (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    if (device < 0) NRB_CUDA(cudaGetDevice(&device));
    if (device >= count) return fail(NRB_ERR_INVALID, "device index out of range");
    NRB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop{};
    NRB_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(NRB_ERR_CUDA, "libnrb200 is built for sm_100a only; device is sm_" + std::to_string(prop.major) +
                                      std::to_string(prop.minor));
    nrb_renderer* r = new nrb_renderer();
    r->device = device;
    int rc = configure_network(r, cfg);
    if (rc != NRB_OK) { delete r; return rc; }
    if (cudaStreamCreateWithFlags(&r->early, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreateWithFlags(&r->evSplit, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&r->evTouch, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&r->evFrameDone[0], cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&r->evFrameDone[1], cudaEventDisableTiming) != cudaSuccess ||
        cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&r->evStart) != cudaSuccess || cudaEventCreate(&r->evStop) != cudaSuccess ||
        cudaEventCreateWithFlags(&r->evJoin, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&r->evJoinIn, cudaEventDisableTiming) != cudaSuccess) {
        delete r;
        return fail(NRB_ERR_CUDA, "stream/event creation failed");
    }
    if (cudaHostAlloc(reinterpret_cast<void**>(&r->pathFeedback), 4 * sizeof(int), cudaHostAllocMapped) == cudaSuccess) {
        r->pathFeedback[0] = r->pathFeedback[1] = -1;
    } else {
        r->pathFeedback = nullptr;
        cudaGetLastError();
    }
    // function attributes are per device: opt every > 48 KB kernel in on THIS device (renderers may live on several GPUs)
    init_frame_kernels();
    init_query_kernels();
    if (cudaGetLastError() != cudaSuccess) {
        nrb_destroy(r);
        return fail(NRB_ERR_CUDA, "kernel attribute setup failed");
    }
    // Material::defaultMaterial(), src/material.h:86-106
    MaterialDev& g = r->globalMaterial;
    std::memset(&g, 0, sizeof(g));
    g.base_color.value[0] = g.base_color.value[1] = g.base_color.value[2] = 0.8f;
    g.clearcoat_gloss.value = 1.0f;
    g.emission_scale = 1.0f;
    g.ior = 1.5f;
    {
        // the reference zero-fills its parameter block at construction and renders with it until a checkpoint is
        // loaded (:1691-1697, neuralReady needs only the network and an outer shell, :2258): same here
        std::vector<uint16_t> zeros(r->gridParams + r->mlpParams, 0);
        rc = nrb_load_weights(r, zeros.data(), zeros.size() * 2, 1);
        if (rc != NRB_OK) { nrb_destroy(r); return rc; }
    }
    *out = r;
    return NRB_OK;
}

void nrb_destroy(nrb_renderer* r) {
    if (!r) return;
    cudaSetDevice(r->device);
    for (nrb_renderer* k : r->kids) nrb_destroy(k);
    r->kids.clear();
    if (r->evFork) cudaEventDestroy(r->evFork);
    if (r->evDone) cudaEventDestroy(r->evDone);
    if (r->early) cudaStreamSynchronize(r->early);
    if (r->stream) cudaStreamSynchronize(r->stream);
    if (r->early) cudaStreamDestroy(r->early);
    if (r->evSplit) cudaEventDestroy(r->evSplit);
    if (r->evTouch) cudaEventDestroy(r->evTouch);
    for (cudaEvent_t e : r->evFrameDone)
        if (e) cudaEventDestroy(e);
    if (r->evStart) cudaEventDestroy(r->evStart);
    if (r->evStop) cudaEventDestroy(r->evStop);
    if (r->evJoin) cudaEventDestroy(r->evJoin);
    if (r->evJoinIn) cudaEventDestroy(r->evJoinIn);
    if (r->gather.remotePixels) cudaIpcCloseMemHandle(r->gather.remotePixels);
    if (r->gather.remoteSync) cudaIpcCloseMemHandle(r->gather.remoteSync);
    if (r->copyStream) { cudaStreamSynchronize(r->copyStream); cudaStreamDestroy(r->copyStream); }
    if (r->evStaged) cudaEventDestroy(r->evStaged);
    if (r->evCopied) cudaEventDestroy(r->evCopied);
    if (r->prof.ready)
        for (int i = 0; i <= StageProfiler::kMaxMarks; ++i) cudaEventDestroy(r->prof.ev[i]);
    if (r->stream) cudaStreamDestroy(r->stream);
    if (r->pathFeedback) cudaFreeHost(r->pathFeedback);
    delete r;
}

int nrb_update_mesh(nrb_renderer* r, int slot, const nrb_mesh_desc* d) {
    if (!r || slot < 0 || slot >= NRB_MESH_SLOT_COUNT) return fail(NRB_ERR_INVALID, "bad renderer/slot");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    MeshSlot& m = r->mesh[slot];
    m.clear();
    if (!d || d->num_triangles <= 0) return NRB_OK;
    // normal cones on every mesh: the shells' facing-mode traces, and the ground-truth primary pass (FORWARD_ONLY on the
    // classic mesh: k_gt_primary 0.364 -> 0.324 ms at 1080p - silhouette rays skip the back-facing leaves they graze)
    int rc = relayout_mesh(r, m, d, true);
    if (rc != NRB_OK) { m.clear(); return rc; }
    MeshDev& v = m.view;
    cudaStream_t s = r->stream;
    NRB_CUDA(m.vertices.upload(d->vertices, sizeof(float) * 3 * d->num_vertices, s));
    NRB_CUDA(m.indices.upload(d->indices, sizeof(uint32_t) * 3 * d->num_triangles, s));
    v.vertices = m.vertices.as<float>();
    v.indices = m.indices.as<uint32_t>();
    if (d->normals) {
        NRB_CUDA(m.normals.upload(d->normals, sizeof(float) * 3 * d->num_vertices, s));
        v.normals = m.normals.as<float>();
    }
    if (d->texcoords) {
        NRB_CUDA(m.texcoords.upload(d->texcoords, sizeof(float) * 2 * d->num_vertices, s));
        v.texcoords = m.texcoords.as<float>();
    }
    if (d->materials && d->num_materials > 0) {
        NRB_CUDA(m.materials.upload(d->materials, sizeof(nrb_material) * d->num_materials, s));
        v.materials = m.materials.as<MaterialDev>();
        v.numMaterials = d->num_materials;
    }
    if (d->material_map && d->material_ids && d->num_primitives > 0) {
        NRB_CUDA(m.materialMap.upload(d->material_map, sizeof(uint32_t) * d->num_primitives, s));
        NRB_CUDA(m.materialIds.upload(d->material_ids, sizeof(int32_t) * d->num_primitives, s));
        v.materialMap = m.materialMap.as<uint32_t>();
        v.materialIds = m.materialIds.as<int>();
        v.numPrimitives = d->num_primitives;
    }
    if (d->textures && d->num_textures > 0) {
        std::vector<TexDev> tv(d->num_textures);
        m.texPixels.resize(d->num_textures);
        for (int i = 0; i < d->num_textures; ++i) {
            const nrb_texture& t = d->textures[i];
            const size_t bytes = static_cast<size_t>(t.width) * t.height * t.channels;
            if (t.pixels && bytes) NRB_CUDA(m.texPixels[i].upload(t.pixels, bytes, s));
            tv[i] = TexDev{m.texPixels[i].as<unsigned char>(), t.width, t.height, t.channels, t.srgb};
        }
        NRB_CUDA(m.textures.upload(tv.data(), sizeof(TexDev) * tv.size(), s));
        v.textures = m.textures.as<TexDev>();
        v.numTextures = d->num_textures;
    }
    v.textureNearest = d->texture_nearest;
    v.hasMeshMaterials = d->has_mesh_materials;
    std::memcpy(m.boundsMin, d->bounds_min, 12);
    std::memcpy(m.boundsMax, d->bounds_max, 12);
    if (slot == NRB_MESH_OUTER_SHELL) m.closed = closed_surface(d->vertices, d->num_vertices, d->indices, d->num_triangles);
    return NRB_OK;
}

int nrb_update_env(nrb_renderer* r, const float* rgb, int width, int height, float strength) {
    if (!r) return fail(NRB_ERR_INVALID, "null renderer");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    r->envPixels.release();
    r->envW = r->envH = 0;
    r->envStrength = strength;
    if (rgb && width > 0 && height > 0) {
        NRB_CUDA(r->envPixels.upload(rgb, sizeof(float) * 3 * static_cast<size_t>(width) * height, r->stream));
        r->envW = width;
        r->envH = height;
    }
    return NRB_OK;
}

int nrb_set_global_material(nrb_renderer* r, const nrb_material* m) {
    if (!r || !m) return fail(NRB_ERR_INVALID, "null argument");
    std::memcpy(&r->globalMaterial, m, sizeof(MaterialDev));
    return NRB_OK;
}

size_t nrb_params_bytes(const nrb_renderer* r) { return (r->gridParams * r->cfg.n_points + r->mlpParams) * 2; }
size_t nrb_weights_file_bytes(const nrb_renderer* r) { return (r->gridParams + r->mlpParams) * 2; }

int nrb_load_weights(nrb_renderer* r, const void* fp16, size_t bytes, int swap_order) {
    if (!r || !fp16) return fail(NRB_ERR_INVALID, "null argument");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    const size_t expect = nrb_weights_file_bytes(r);
    if (bytes != expect)
        return fail(NRB_ERR_SIZE, "Weights file size mismatch: expected " + std::to_string(expect) + " bytes, got " +
                                      std::to_string(bytes));
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    const uint16_t* file = static_cast<const uint16_t*>(fp16);
    const uint16_t* srcMlp = swap_order ? file + r->gridParams : file;
    const uint16_t* srcGrid = swap_order ? file : file + r->mlpParams;
    NRB_CUDA(r->grid.upload(srcGrid, r->gridParams * 2, r->stream));
    NRB_CUDA(r->mlp.upload(srcMlp, r->mlpParams * 2, r->stream));
    r->net.grid = r->grid.as<__half>();
    r->net.mlp = r->mlp.as<__half>();
    r->net.packedW = nullptr;
    if (split_query_supported(r->net)) {  // UMMA-operand image of the layers for k_query_mlp
        NRB_CUDA(r->packedW.alloc(split_packed_bytes(r->net)));
        r->net.packedW = r->packedW.p;
        if (split_prepack(r->net, r->packedW.p, r->stream) != 0) return fail(NRB_ERR_CUDA, "weight pre-pack launch failed");
    }
    r->weightsLoaded = true;
    return NRB_OK;
}

int nrb_load_weights_file(nrb_renderer* r, const char* path, int swap_order) {
    if (!r || !path) return fail(NRB_ERR_INVALID, "null argument");
    FILE* f = std::fopen(path, "rb");
    if (!f) return fail(NRB_ERR_INVALID, std::string("Failed to open weights file: ") + path);
    std::fseek(f, 0, SEEK_END);
    const long size = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<unsigned char> data(size > 0 ? static_cast<size_t>(size) : 0);
    const size_t got = data.empty() ? 0 : std::fread(data.data(), 1, data.size(), f);
    std::fclose(f);
    if (got != data.size()) return fail(NRB_ERR_INVALID, std::string("Failed to read weights file: ") + path);
    return nrb_load_weights(r, data.data(), data.size(), swap_order);
}

int nrb_resize(nrb_renderer* r, int width, int height) {
    if (!r) return fail(NRB_ERR_INVALID, "null renderer");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    if (width == r->width && height == r->height) return NRB_OK;
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    r->width = width;
    r->height = height;
    // RendererNeural::resize frees every buffer (src/cuda_renderer_neural.cu:1736-1746)
    r->pool.clear();
    r->accum.release();
    r->pixels.release();
    r->stateElems = r->accumPixels = r->pixelCount = 0;
    r->accumSampleCount = 0;
    gather_reset(r);  // the exported frame buffer is gone: export / attach again
    return NRB_OK;
}

int nrb_set_camera_basis(nrb_renderer* r, const float forward[3], const float right[3], const float up[3], float fov_y) {
    if (!r) return fail(NRB_ERR_INVALID, "null renderer");
    std::memcpy(r->fwd, forward, 12);
    std::memcpy(r->right, right, 12);
    std::memcpy(r->up, up, 12);
    r->fovY = fov_y;
    return NRB_OK;
}

int nrb_set_option(nrb_renderer* r, int option, int value) {
    if (!r) return fail(NRB_ERR_INVALID, "null renderer");
    switch (option) {
        case NRB_OPT_USE_NEURAL_QUERY: r->useNeuralQuery = value != 0; break;
        case NRB_OPT_SAMPLES_PER_PIXEL: r->samplesPerPixel = value; break;
        case NRB_OPT_BOUNCE_COUNT: r->bounceCount = value; break;
        case NRB_OPT_LAMBERT_VIEW: r->lambertView = value != 0; break;
        case NRB_OPT_CLASSIC_MESH_INDEX: r->classicMeshIndex = value; break;
        case NRB_OPT_USE_ADDITIONAL_MESH: r->useAdditionalMesh = value != 0; break;
        case NRB_OPT_USE_CONSTANT_NEURAL_COLOR: r->useConstantNeuralColor = value != 0; break;
        case NRB_OPT_USE_DIRECT_ENV_COLOR: r->useDirectEnvColor = value != 0; break;
        case NRB_OPT_CAPTURE: r->capture = value; break;
        case NRB_OPT_PROFILE: r->profile = value; break;
        case NRB_OPT_QUERY_KERNEL: r->net.kernelChoice = value; break;
        case NRB_OPT_TAIL_GRAPH: r->tailMode = value < 0 ? 0 : (value > 2 ? 2 : value); break;
        case NRB_OPT_DENOISE: r->denoise = value != 0; break;
        case NRB_OPT_LITERAL_BOUNCE_TRACES: r->literalBounceTraces = value != 0; break;
        case NRB_OPT_BOUNCE_ENTRY: r->bounceEntry = value != 0; break;
        case NRB_OPT_TILE_ORDER: r->tileOrderOn = value != 0; r->tileOrderState = 0; break;
        case NRB_OPT_FRAME_SLICES: r->slices = value < 0 ? 0 : (value > 8 ? 8 : value); break;
        case NRB_OPT_PIPELINE: r->pipeline = value != 0; break;
        case NRB_OPT_PATH_TAIL:
            r->pathTailBelow = value < 0 ? 0 : value;
            if (r->pathFeedback) r->pathFeedback[0] = r->pathFeedback[1] = -1;
            break;
        case NRB_OPT_USE_HARDWARE_RT: break;  // accepted, ignored: there is no OptiX path
        default: return fail(NRB_ERR_INVALID, "unknown option");
    }
    return NRB_OK;
}

int nrb_get_option(const nrb_renderer* r, int option) {
    switch (option) {
        case NRB_OPT_USE_NEURAL_QUERY: return r->useNeuralQuery;
        case NRB_OPT_SAMPLES_PER_PIXEL: return r->samplesPerPixel;
        case NRB_OPT_BOUNCE_COUNT: return r->bounceCount;
        case NRB_OPT_LAMBERT_VIEW: return r->lambertView;
        case NRB_OPT_CLASSIC_MESH_INDEX: return r->classicMeshIndex;
        case NRB_OPT_USE_ADDITIONAL_MESH: return r->useAdditionalMesh;
        case NRB_OPT_USE_CONSTANT_NEURAL_COLOR: return r->useConstantNeuralColor;
        case NRB_OPT_USE_DIRECT_ENV_COLOR: return r->useDirectEnvColor;
        case NRB_OPT_CAPTURE: return r->capture;
        case NRB_OPT_PROFILE: return r->profile;
        case NRB_OPT_QUERY_KERNEL: return r->net.kernelChoice;
        case NRB_OPT_TAIL_GRAPH: return r->tailMode;
        case NRB_OPT_DENOISE: return r->denoise ? 1 : 0;
        case NRB_OPT_LITERAL_BOUNCE_TRACES: return r->literalBounceTraces ? 1 : 0;
        case NRB_OPT_BOUNCE_ENTRY: return r->bounceEntry;
        case NRB_OPT_PATH_TAIL: return r->pathTailBelow;
        case NRB_OPT_TILE_ORDER: return r->tileOrderOn ? 1 : 0;
        case NRB_OPT_FRAME_SLICES: return r->slices;
        case NRB_OPT_PIPELINE: return r->pipeline;
        default: return 0;
    }
}

int nrb_set_envmap_rotation(nrb_renderer* r, float degrees) { r->envmapRotation = degrees; return NRB_OK; }
int nrb_set_constant_neural_color(nrb_renderer* r, const float rgb[3]) { std::memcpy(r->constantNeuralColor, rgb, 12); return NRB_OK; }
int nrb_set_direct_env_color(nrb_renderer* r, const float rgb[3]) { std::memcpy(r->directEnvColor, rgb, 12); return NRB_OK; }

int nrb_set_tile(nrb_renderer* r, int band_rows, int rank, int world) {
    if (!r || band_rows <= 0 || world <= 0 || rank < 0 || rank >= world) return fail(NRB_ERR_INVALID, "bad tile spec");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    if (band_rows != r->bandRows || rank != r->rank || world != r->world) {
        r->bandRows = band_rows; r->rank = rank; r->world = world;
        r->pool.clear();
        r->accum.release();
        r->stateElems = r->accumPixels = 0;
        r->accumSampleCount = 0;
    }
    return NRB_OK;
}

int nrb_width(const nrb_renderer* r) { return r->width; }
int nrb_height(const nrb_renderer* r) { return r->height; }
uint32_t nrb_accum_sample_count(const nrb_renderer* r) { return r->accumSampleCount; }
void* nrb_device_pixels(const nrb_renderer* r) { return r->pixels.p; }
void* nrb_stream(const nrb_renderer* r) { return r->stream; }
int nrb_input_dims(const nrb_renderer* r) { return r->net.inputDims; }
int nrb_output_dims(const nrb_renderer*) { return 16; }

int nrb_sync(nrb_renderer* r) {
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    if (r->copyStream) NRB_CUDA(cudaStreamSynchronize(r->copyStream));
    r->copyPending = false;
    if (r->gather.role != 0) {
        unsigned err = 0;
        NRB_CUDA(cudaMemcpy(&err, r->gather.sync.as<unsigned>() + 4, sizeof(err), cudaMemcpyDeviceToHost));
        if (err) return fail(NRB_ERR_CUDA, "frame gather timed out: a rank stopped rendering (frames are out of step)");
    }
    return NRB_OK;
}

}  // extern "C"

// =================================================================================================
// Buffers and the frame
// =================================================================================================
namespace {

// ensureNetworkBuffers / ensureAccumBuffer (src/cuda_renderer_neural.cu:2826-2975): one pool of
// 16-byte-record SoA arrays sized for the elements this renderer owns.
int ensure_state(nrb_renderer* r, size_t nElems, size_t localPixels, size_t framePixels, int buffers) {
    if (nElems != r->stateElems) {
        if (r->early) NRB_CUDA(cudaStreamSynchronize(r->early));
        if (r->stream) NRB_CUDA(cudaStreamSynchronize(r->stream));
        r->pool.clear();
        r->st = State{};
        r->stBuf[0] = r->stBuf[1] = State{};
        r->stBufValid[0] = r->stBufValid[1] = false;
        r->frameDoneValid[0] = r->frameDoneValid[1] = false;
        r->stateElems = nElems;
    }
    for (int b = 0; b < buffers; ++b) {
        if (r->stBufValid[b]) continue;
        const size_t n = (nElems + 255) / 256 * 256;
        auto take = [&](size_t bytes, void** out) -> cudaError_t {
            r->pool.emplace_back();
            cudaError_t e = r->pool.back().alloc(bytes);
            *out = r->pool.back().p;
            return e;
        };
        State& s = r->stBuf[b];
        for (int k = 0; k < 3; ++k) {
            NRB_CUDA(take(n * 16, reinterpret_cast<void**>(&s.rec[k].P)));
            NRB_CUDA(take(n * 16, reinterpret_cast<void**>(&s.rec[k].N)));
            NRB_CUDA(take(n * 16, reinterpret_cast<void**>(&s.rec[k].C)));
            NRB_CUDA(take(n * 8, reinterpret_cast<void**>(&s.rec[k].M)));
            if (k == 1) break;  // rec[2] is reserved (the hybrid-mesh hit is merged in registers)
        }
        float4** f4s[] = {&s.T, &s.L, &s.rayO, &s.rayD, &s.rayW, &s.prevD, &s.segE, &s.segD, &s.segX};
        for (float4** p : f4s) NRB_CUDA(take(n * 16, reinterpret_cast<void**>(p)));
        NRB_CUDA(take(n * 8, reinterpret_cast<void**>(&s.segI)));
        for (int q = 0; q < 4; ++q) NRB_CUDA(take(n * 4, reinterpret_cast<void**>(&s.queue[q])));
        NRB_CUDA(take(64 * sizeof(int), reinterpret_cast<void**>(&s.counts)));
        NRB_CUDA(take(16 * sizeof(unsigned long long), reinterpret_cast<void**>(&s.stats)));
        NRB_CUDA(take(n * r->net.inputDims * sizeof(float), reinterpret_cast<void**>(&s.netIn)));
        NRB_CUDA(take(n * 16 * sizeof(__half), reinterpret_cast<void**>(&s.netOut)));
        r->stBufValid[b] = true;
    }
    if (localPixels != r->accumPixels) {
        NRB_CUDA(r->accum.alloc(localPixels * 3 * sizeof(float)));
        NRB_CUDA(cudaMemsetAsync(r->accum.p, 0, r->accum.bytes, r->stream));
        r->accumPixels = localPixels;
        r->accumSampleCount = 0;
    }
    if (!r->parent && framePixels != r->pixelCount) {
        NRB_CUDA(r->pixels.alloc(framePixels * sizeof(uchar4)));
        NRB_CUDA(cudaMemsetAsync(r->pixels.p, 0, r->pixels.bytes, r->stream));
        r->pixelCount = framePixels;
    }
    {
        const int tiles = static_cast<int>(((r->width + 7) / 8) * ((r->localHeight + 3) / 4)) * (nElems && localPixels ? static_cast<int>(nElems / localPixels) : 1);
        if (tiles != r->tileCount) {
            NRB_CUDA(r->tileCost.alloc(sizeof(unsigned int) * tiles));
            NRB_CUDA(r->tileCostSorted.alloc(sizeof(unsigned int) * tiles));
            NRB_CUDA(r->tileIds.alloc(sizeof(int) * tiles));
            NRB_CUDA(r->tileOrder.alloc(sizeof(int) * tiles));
            NRB_CUDA(r->tileSortTemp.alloc(tile_sort_temp_bytes(tiles)));
            r->tileCount = tiles;
            r->tileOrderState = 0;
        }
    }
    for (int b = 0; b < 2; ++b) {
        r->stBuf[b].accum = r->accum.as<float>();
        r->stBuf[b].pixels = r->parent ? r->pixelsOverride
                                       : (r->gather.role == 2 ? static_cast<uchar4*>(r->gather.remotePixels) : r->pixels.as<uchar4>());
    }
    if (!r->stBufValid[r->parity]) r->parity = 0;
    r->st = r->stBuf[r->parity];
    return NRB_OK;
}

// Scratch for the split query's encoded tiles: at most 2^22 rows (the launcher walks larger batches
// in chunks), grown on demand.
int ensure_xt(nrb_renderer* r, size_t rows) {
    if (!r->net.packedW || !split_query_supported(r->net)) {
        r->net.xt = nullptr;
        r->net.xtRows = 0;
        return NRB_OK;
    }
    size_t cap = (rows + 127) / 128 * 128;
    if (cap > (1u << 22)) cap = 1u << 22;
    if (cap < 128) cap = 128;
    const size_t need = split_xt_bytes(r->net, cap);
    if (r->xt.bytes < need) {
        NRB_CUDA(cudaStreamSynchronize(r->stream));
        NRB_CUDA(r->xt.alloc(need));
    }
    if (!r->encCur.p) {
        NRB_CUDA(r->encCur.alloc(kMaxLevels * sizeof(int)));
        NRB_CUDA(cudaMemsetAsync(r->encCur.p, 0, r->encCur.bytes, r->stream));
    }
    r->net.encCursor = r->encCur.as<int>();
    r->net.xt = r->xt.p;
    r->net.xtRows = static_cast<uint32_t>(r->xt.bytes / split_xt_bytes(r->net, 128) * 128);
    return NRB_OK;
}

// The reference's KernelId bucket (src/cuda_renderer_neural.h:17-34) a kernel of this library reports into.  Stages
// are fused across the reference's buckets, so a kernel goes to the bucket of its dominant part (named in the list).
int timing_bucket(const std::string& k) {
    if (k == "k_gt_primary") return 0;                                   // KID_primaryTrace (+ init path state)
    if (k == "k_gt_bounce") return 1;                                    // KID_bounceTrace (+ sample, integrate)
    if (k == "k_bounce_begin") return 8;                                 // KID_earlyTermination (+ the shell entry of bounce rays)
    if (k.rfind("k_query_", 0) == 0 || k.rfind("k_neural_query", 0) == 0) return 3;  // KID_neuralForward
    if (k == "k_primary_finish") return 4;                               // KID_additionalMeshPrimary (+ select closest, init path state)
    if (k == "k_bounce_sample") return 7;                                // KID_sampleBounce
    if (k == "k_bounce_finish") return 11;                               // KID_integrateBounce (+ additional mesh, select closest)
    if (k == "k_finalize" || k == "k_lambert" || k == "k_denoise" || k.rfind("k_gather", 0) == 0) return 12;  // KID_finalize
    return 2;                                                            // KID_shellTrace: entry, exits, apply, segment tail
}

// Turn the event marks of the last profiled frame into per-kernel totals.
int drain_profiler(nrb_renderer* r) {
    if (!r->profPending) return NRB_OK;
    r->profPending = false;
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    for (double& b : r->lastBuckets) b = 0.0;
    r->lastBucketsValid = r->prof.marks > 0;
    for (int i = 0; i < r->prof.marks; ++i) {
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, r->prof.ev[i], r->prof.ev[i + 1]) != cudaSuccess) continue;
        const std::string key(r->prof.name[i]);
        r->lastBuckets[timing_bucket(key)] += ms;
        auto it = r->stages.find(key);
        if (it == r->stages.end()) {
            r->stageOrder.push_back(key);
            it = r->stages.emplace(key, std::make_pair(0.0, 0ull)).first;
        }
        it->second.first += ms;
        it->second.second += 1;
    }
    r->prof.marks = 0;
    return NRB_OK;
}

bool differs(const float* a, const float* b, float eps) {
    return std::fabs(a[0] - b[0]) > eps || std::fabs(a[1] - b[1]) > eps || std::fabs(a[2] - b[2]) > eps;
}

V3 to_v3(const float* p) { return V3{p[0], p[1], p[2]}; }

}  // namespace

extern "C" {

int nrb_reset_samples(nrb_renderer* r) {
    if (!r) return fail(NRB_ERR_INVALID, "null renderer");
    NRB_CUDA(cudaSetDevice(r->device));
    for (nrb_renderer* k : r->kids) {
        if (k->accum.p) NRB_CUDA(cudaMemsetAsync(k->accum.p, 0, k->accum.bytes, k->stream));
        k->accumSampleCount = 0;
    }
    if (r->accum.p) NRB_CUDA(cudaMemsetAsync(r->accum.p, 0, r->accum.bytes, r->stream));
    r->accumSampleCount = 0;
    return NRB_OK;
}

static int render_impl(nrb_renderer* r, const float cam_pos[3]);
static int render_sliced(nrb_renderer* r, const float cam_pos[3]);

int nrb_render(nrb_renderer* r, const float cam_pos[3]) {
    if (!r || !cam_pos) return fail(NRB_ERR_INVALID, "null argument");
    if (r->width <= 0 || r->height <= 0) return NRB_OK;  // the reference returns silently (:2086-2091)
    NRB_CUDA(cudaSetDevice(r->device));
    // slices need the frame's stages to be plain launches on one stream each: not with the per-launch profiler, the
    // debug capture buffers (whole-frame element order), the denoiser (whole-frame window) or the graph-driven tail
    // Measured at 1080p: ground truth 1.25 -> 1.06 ms with 4 slices (6.03 -> 4.98 ms at 8 spp) - its
    // traversal launches end on a few long silhouette rays, which the other slices' work covers; the neural path
    // gets slower (1.39 -> 1.43 ms with 2): its stages already fill the machine and the query kernels of two slices
    // cannot share an SM's shared memory.  Hence the automatic choice.
    const bool neuralActive = r->useNeuralQuery && r->weightsLoaded && r->mesh[NRB_MESH_OUTER_SHELL].view.numTriangles > 0;
    // ground truth: pipelined unsliced frames measured best (0.759 ms against 0.820 with 4 slices and 0.848 with neither)
    const bool canPipeline = r->pipeline && !r->lambertView && !r->profile && !r->capture && !r->denoise && !debug_sync_enabled();
    const int S = r->slices > 0 ? r->slices : ((neuralActive || canPipeline) ? 1 : 4);
    const bool sliced = S > 1 && !r->profile && !r->capture && !r->denoise && !debug_sync_enabled() &&
                        (r->tailMode == 2 || !neuralActive) && local_rows(r) >= 64 * S;
    r->activeSlices = sliced ? S : 1;
    if (sliced != r->lastSliced) {  // the accumulation lives in different buffers: start over
        int rc = nrb_reset_samples(r);
        if (rc != NRB_OK) return rc;
        r->hasLastCamera = false;
        for (nrb_renderer* k : r->kids) k->hasLastCamera = false;
        r->lastSliced = sliced;
    }
    if (sliced) r->touched = true;  // the slices' streams join `stream`: a pipelined frame after them waits for it
    return sliced ? render_sliced(r, cam_pos) : render_impl(r, cam_pos);
}

static int render_impl(nrb_renderer* r, const float cam_pos[3]) {
    const int maxBounces = std::max(0, r->bounceCount);
    const int spp = r->samplesPerPixel <= 0 ? 1 : r->samplesPerPixel;

    // classic mesh selection (:2119-2125)
    int classicSlot = NRB_MESH_ORIGINAL;
    if (r->classicMeshIndex == 1 && r->mesh[NRB_MESH_INNER_SHELL].view.numTriangles > 0) classicSlot = NRB_MESH_INNER_SHELL;
    else if (r->classicMeshIndex == 2 && r->mesh[NRB_MESH_OUTER_SHELL].view.numTriangles > 0) classicSlot = NRB_MESH_OUTER_SHELL;

    if (r->denoise && r->world > 1) return fail(NRB_ERR_INVALID, "NRB_OPT_DENOISE needs the whole frame on one GPU (nrb_set_tile world 1)");
    const size_t framePixels = static_cast<size_t>(r->width) * r->height;
    r->localHeight = local_rows(r);
    const size_t localPixels = static_cast<size_t>(r->width) * r->localHeight;
    const size_t nElems = localPixels * spp;
    if (nElems >= (1ull << 31)) return fail(NRB_ERR_INVALID, "width*height*spp exceeds 2^31 elements; render in smaller spp batches");
    const MeshSlot& outer = r->mesh[NRB_MESH_OUTER_SHELL];
    const bool neural = r->useNeuralQuery && r->weightsLoaded && outer.view.numTriangles > 0;  // (:2258-2260)
    // pipelined frames: see nrb_renderer::pipeline.  Plain launches on two streams only - not with the per-launch
    // profiler, the capture buffers, the graph-driven tail or as a slice.
    const bool pipelined = r->pipeline && !r->parent && !r->lambertView && !r->profile && !r->capture && !r->denoise &&
                           (r->tailMode == 2 || !neural) && !debug_sync_enabled() && r->early != nullptr;
    int rc = ensure_state(r, nElems, localPixels, framePixels, pipelined ? 2 : 1);
    if (rc != NRB_OK) return rc;
    if (r->useNeuralQuery && r->weightsLoaded) {
        rc = ensure_xt(r, nElems);
        if (rc != NRB_OK) return rc;
    }

    // accumulation reset rules (:2190-2217)
    bool cameraMoved = !r->hasLastCamera;
    const float kEps = 1e-4f;
    if (!cameraMoved)
        cameraMoved = differs(cam_pos, r->lastCam, kEps) || differs(r->fwd, r->lastFwd, kEps) ||
                      differs(r->right, r->lastRight, kEps) || differs(r->up, r->lastUp, kEps) ||
                      std::fabs(r->fovY - r->lastFovY) > kEps;
    if (cameraMoved || r->useNeuralQuery != r->lastUseNeuralQuery || r->lambertView != r->lastLambertView ||
        maxBounces != r->lastBounceCount || spp != r->lastSamplesPerPixel || r->classicMeshIndex != r->lastClassicMeshIndex ||
        r->envmapRotation != r->lastEnvmapRotation || r->useAdditionalMesh != r->lastUseAdditionalMesh) {
        rc = nrb_reset_samples(r);
        if (rc != NRB_OK) return rc;
        if (r->tileOrderState == 2) r->tileOrderState = 3;  // keep the old order (still a permutation), re-sort after this frame
    }
    r->lastUseNeuralQuery = r->useNeuralQuery; r->lastLambertView = r->lambertView; r->lastBounceCount = maxBounces;
    r->lastSamplesPerPixel = spp; r->lastClassicMeshIndex = r->classicMeshIndex; r->lastEnvmapRotation = r->envmapRotation;
    r->lastUseAdditionalMesh = r->useAdditionalMesh; r->lastFovY = r->fovY; r->hasLastCamera = true;
    std::memcpy(r->lastCam, cam_pos, 12); std::memcpy(r->lastFwd, r->fwd, 12);
    std::memcpy(r->lastRight, r->right, 12); std::memcpy(r->lastUp, r->up, 12);

    Frame f{};
    FrameParams& fp = f.fp;
    fp.camPos = to_v3(cam_pos); fp.camForward = to_v3(r->fwd); fp.camRight = to_v3(r->right); fp.camUp = to_v3(r->up);
    fp.outerMin = to_v3(outer.boundsMin);
    for (int k = 0; k < 3; ++k) {  // (:2171-2178)
        const float ext = outer.boundsMax[k] - outer.boundsMin[k];
        (&fp.outerInvExtent.x)[k] = ext != 0.0f ? 1.0f / ext : 0.0f;
    }
    const MeshSlot& classic = r->mesh[classicSlot];
    const float sx = classic.boundsMax[0] - classic.boundsMin[0], sy = classic.boundsMax[1] - classic.boundsMin[1],
                sz = classic.boundsMax[2] - classic.boundsMin[2];
    float sceneScale = sqrtf(sx * sx + sy * sy + sz * sz);  // (:2180-2189)
    if (sceneScale < 1e-6f) sceneScale = 1.0f;
    fp.constantNeuralColor = to_v3(r->constantNeuralColor);
    fp.directEnvColor = to_v3(r->directEnvColor);
    fp.tanHalfFov = tanf(r->fovY * 0.5f);
    fp.aspect = static_cast<float>(r->width) / static_cast<float>(r->height);
    fp.maxRadiance = 100.0f;
    fp.sceneScale = sceneScale;
    fp.useConstantNeuralColor = r->useConstantNeuralColor;
    fp.useDirectEnvColor = r->useDirectEnvColor;
    fp.maxBounces = maxBounces;
    fp.width = r->width; fp.height = r->height; fp.pixelCount = static_cast<int>(framePixels); fp.spp = spp;
    fp.sampleOffset = r->accumSampleCount;
    fp.bandRows = r->bandRows; fp.rank = r->rank; fp.world = r->world; fp.localHeight = r->localHeight;
    fp.localPixels = static_cast<int>(localPixels);
    fp.nElems = static_cast<int>(nElems);
    fp.nPoints = r->net.nPoints; fp.inputDims = r->net.inputDims;
    for (int k = 0; k < 4; ++k) f.mesh[k] = r->mesh[k].view;
    if (!r->useAdditionalMesh) f.mesh[NRB_MESH_ADDITIONAL] = MeshDev{};
    f.env = EnvDev{r->envPixels.as<float>(), r->envW, r->envH, r->envmapRotation, r->envStrength};
    f.gmat = r->globalMaterial;
    if (pipelined) {
        r->parity ^= 1;
        r->st = r->stBuf[r->parity];
    }
    f.st = r->st;
    f.st.tileOrder = nullptr;
    f.st.tileCost = nullptr;
    f.classicSlot = classicSlot;
    f.additionalEnabled = r->useAdditionalMesh ? 1 : 0;
    f.literalBounceTraces = r->literalBounceTraces ? 1 : 0;
    f.bounceEntryVote = r->bounceEntry;
    f.literalSegExits = (r->literalBounceTraces || !outer.closed) ? 1 : 0;

    // sOut: the stream the frame completes on.  s: where it starts (the same stream unless pipelined).
    cudaStream_t sOut = r->stream;
    cudaStream_t s = pipelined ? r->early : r->stream;
    if (pipelined) {
        if (r->touched) {  // scene / weight uploads, batch queries or an unpipelined frame since the last frame: wait for them
            NRB_CUDA(cudaEventRecord(r->evTouch, sOut));
            NRB_CUDA(cudaStreamWaitEvent(s, r->evTouch, 0));
        }
        // this state buffer is free once the frame before the last has completed
        if (r->frameDoneValid[r->parity]) NRB_CUDA(cudaStreamWaitEvent(s, r->evFrameDone[r->parity], 0));
    }
    r->touched = !pipelined;
    if (r->profile) {
        rc = drain_profiler(r);
        if (rc != NRB_OK) return rc;
        if (!r->prof.ready) {
            for (int i = 0; i <= StageProfiler::kMaxMarks; ++i) NRB_CUDA(cudaEventCreate(&r->prof.ev[i]));
            r->prof.ready = true;
        }
    }
    // symmetric join: whatever the caller enqueued on the legacy default stream since the last frame (an async copy or
    // a kernel reading devicePixels()) finishes before this frame may overwrite the buffers
    if (!r->parent) {
        NRB_CUDA(cudaEventRecord(r->evJoinIn, cudaStreamLegacy));
        NRB_CUDA(cudaStreamWaitEvent(sOut, r->evJoinIn, 0));  // only the end of the frame (finalize) writes them
    }
    NRB_CUDA(cudaMemsetAsync(r->st.stats, 0, 16 * sizeof(unsigned long long), s));
    NRB_CUDA(cudaEventRecord(r->evStart, s));
    if (r->parent) {
        // a slice: the parent does the hand-shakes of the frame gather and the stream joins
    } else if (r->gather.role == 1) {
        // everything the caller enqueued on this stream since the last frame has read it: peers may overwrite
        launch_gather_release(r->gather.sync.as<unsigned>(), r->gather.frames, sOut);
    } else if (r->gather.role == 2) {
        launch_gather_wait(static_cast<unsigned*>(r->gather.remoteSync) + 1, r->gather.frames, r->gather.sync.as<unsigned>() + 4, sOut);
    }
    if (r->profile) {
        r->prof.marks = 0;
        NRB_CUDA(cudaEventRecord(r->prof.ev[0], s));
        set_stage_profiler(&r->prof);
    }
    int launches;
    if (neural) launches = launch_neural_frame(f, r->net, r->lambertView, s, (r->tailMode == 1 && !debug_sync_enabled()) ? &r->segTail : nullptr, r->tailMode == 2,
                                               r->pathFeedback, r->tailMode == 2 ? r->pathTailBelow : 0, sOut, r->evSplit);
    else {
        // primary tiles most-expensive-first, from the clocks an earlier frame of this camera measured (see k_gt_primary)
        const bool order = r->tileOrderOn && r->tileCount > 0;
        if (order) {
            f.st.tileCost = r->tileOrderState == 2 ? nullptr : r->tileCost.as<unsigned int>();
            f.st.tileOrder = r->tileOrderState >= 2 ? r->tileOrder.as<int>() : nullptr;
        }
        launches = launch_gt_frame(f, r->lambertView, s, r->pathFeedback, r->pathTailBelow, sOut, r->evSplit);
        if (order && r->tileOrderState != 2) {
            if (launch_tile_sort(r->tileCost.as<unsigned int>(), r->tileCostSorted.as<unsigned int>(), r->tileIds.as<int>(),
                                 r->tileOrder.as<int>(), r->tileCount, r->tileSortTemp.p, r->tileSortTemp.bytes, s) == 0)
                r->tileOrderState = 2;
            launches += 2;
        }
    }
    if (r->denoise && !r->lambertView) launches += launch_denoise(f, sOut);
    if (!r->parent && r->gather.role != 0) {
        ++r->gather.frames;
        if (r->gather.role == 2) launch_gather_signal(static_cast<unsigned*>(r->gather.remoteSync), sOut);
        else launch_gather_wait(r->gather.sync.as<unsigned>(), r->gather.frames * static_cast<unsigned>(r->gather.peers),
                                r->gather.sync.as<unsigned>() + 4, sOut);
        launches += 2;
    }
    set_stage_profiler(nullptr);
    r->profPending = r->profile != 0;
    NRB_CUDA(cudaEventRecord(r->evStop, sOut));
    if (pipelined) {
        NRB_CUDA(cudaEventRecord(r->evFrameDone[r->parity], sOut));
        r->frameDoneValid[r->parity] = true;
    }
    NRB_CUDA(cudaGetLastError());
    if (r->lambertView) r->accumSampleCount = 0;
    else r->accumSampleCount += static_cast<uint32_t>(spp);
    r->lastNeural = neural;
    r->lastLaunches = launches;
    r->lastRayCount = static_cast<int>(nElems);
    if (r->parent) return NRB_OK;
    // join to the legacy default stream: a following cudaMemcpy(devicePixels) sees the frame
    NRB_CUDA(cudaEventRecord(r->evJoin, sOut));
    NRB_CUDA(cudaStreamWaitEvent(cudaStreamLegacy, r->evJoin, 0));
    return NRB_OK;
}

// One slice renderer: own stream, events and feedback block; everything else arrives through sync_slice.
static nrb_renderer* make_slice(nrb_renderer* r) {
    nrb_renderer* k = new nrb_renderer();
    k->device = r->device;
    k->parent = r;
    if (cudaStreamCreateWithFlags(&k->stream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreate(&k->evStart) != cudaSuccess ||
        cudaEventCreate(&k->evStop) != cudaSuccess || cudaEventCreateWithFlags(&k->evDone, cudaEventDisableTiming) != cudaSuccess) {
        nrb_destroy(k);
        return nullptr;
    }
    if (cudaHostAlloc(reinterpret_cast<void**>(&k->pathFeedback), 4 * sizeof(int), cudaHostAllocMapped) == cudaSuccess) {
        k->pathFeedback[0] = k->pathFeedback[1] = -1;
    } else {
        k->pathFeedback = nullptr;
        cudaGetLastError();
    }
    return k;
}

// Bring a slice up to date with its parent: scene views, weights (borrowed pointers), options, camera, tile.
static int sync_slice(nrb_renderer* r, nrb_renderer* k, int index, uchar4* pixels) {
    for (int m = 0; m < 4; ++m) {
        k->mesh[m].view = r->mesh[m].view;
        std::memcpy(k->mesh[m].boundsMin, r->mesh[m].boundsMin, 12);
        std::memcpy(k->mesh[m].boundsMax, r->mesh[m].boundsMax, 12);
        k->mesh[m].depth = r->mesh[m].depth;
        k->mesh[m].closed = r->mesh[m].closed;
    }
    k->envPixels.alias(r->envPixels.p, r->envPixels.bytes);
    k->envW = r->envW; k->envH = r->envH; k->envStrength = r->envStrength;
    k->globalMaterial = r->globalMaterial;
    k->cfg = r->cfg;
    {
        NetDev n = r->net;  // the slice keeps its own query scratch
        n.xt = k->net.xt; n.xtRows = k->net.xtRows; n.encCursor = k->net.encCursor;
        k->net = n;
    }
    k->gridParams = r->gridParams; k->mlpParams = r->mlpParams; k->weightsLoaded = r->weightsLoaded;
    k->tailMode = r->tailMode; k->literalBounceTraces = r->literalBounceTraces; k->bounceEntry = r->bounceEntry; k->pathTailBelow = r->pathTailBelow;
    k->tileOrderOn = r->tileOrderOn;
    std::memcpy(k->fwd, r->fwd, 12); std::memcpy(k->right, r->right, 12); std::memcpy(k->up, r->up, 12);
    k->fovY = r->fovY;
    k->lambertView = r->lambertView; k->useNeuralQuery = r->useNeuralQuery; k->samplesPerPixel = r->samplesPerPixel;
    k->bounceCount = r->bounceCount; k->classicMeshIndex = r->classicMeshIndex; k->envmapRotation = r->envmapRotation;
    k->useConstantNeuralColor = r->useConstantNeuralColor; k->useDirectEnvColor = r->useDirectEnvColor;
    k->useAdditionalMesh = r->useAdditionalMesh;
    std::memcpy(k->constantNeuralColor, r->constantNeuralColor, 12);
    std::memcpy(k->directEnvColor, r->directEnvColor, 12);
    k->pixelsOverride = pixels;
    int rc = NRB_OK;
    if (k->width != r->width || k->height != r->height) rc = nrb_resize(k, r->width, r->height);
    if (rc == NRB_OK) rc = nrb_set_tile(k, r->bandRows, r->rank + index * r->world, r->world * r->activeSlices);  // bands b with b % world == rank, dealt to the slices
    return rc;
}

static int render_sliced(nrb_renderer* r, const float cam_pos[3]) {
    const int S = r->activeSlices;
    while (static_cast<int>(r->kids.size()) > S) { nrb_destroy(r->kids.back()); r->kids.pop_back(); }
    while (static_cast<int>(r->kids.size()) < S) {
        nrb_renderer* k = make_slice(r);
        if (!k) return fail(NRB_ERR_CUDA, "slice renderer creation failed");
        r->kids.push_back(k);
    }
    if (!r->evFork) NRB_CUDA(cudaEventCreateWithFlags(&r->evFork, cudaEventDisableTiming));
    const size_t framePixels = static_cast<size_t>(r->width) * r->height;
    if (framePixels != r->pixelCount) {
        NRB_CUDA(cudaStreamSynchronize(r->stream));
        NRB_CUDA(r->pixels.alloc(framePixels * sizeof(uchar4)));
        NRB_CUDA(cudaMemsetAsync(r->pixels.p, 0, r->pixels.bytes, r->stream));
        r->pixelCount = framePixels;
    }
    uchar4* target = r->gather.role == 2 ? static_cast<uchar4*>(r->gather.remotePixels) : r->pixels.as<uchar4>();
    cudaStream_t s = r->stream;
    NRB_CUDA(cudaEventRecord(r->evJoinIn, cudaStreamLegacy));
    NRB_CUDA(cudaStreamWaitEvent(s, r->evJoinIn, 0));
    NRB_CUDA(cudaEventRecord(r->evStart, s));
    if (r->gather.role == 1) launch_gather_release(r->gather.sync.as<unsigned>(), r->gather.frames, s);
    else if (r->gather.role == 2)
        launch_gather_wait(static_cast<unsigned*>(r->gather.remoteSync) + 1, r->gather.frames, r->gather.sync.as<unsigned>() + 4, s);
    NRB_CUDA(cudaEventRecord(r->evFork, s));
    int launches = 0, rays = 0;
    for (int i = 0; i < S; ++i) {
        nrb_renderer* k = r->kids[i];
        int rc = sync_slice(r, k, i, target);
        if (rc != NRB_OK) return rc;
        NRB_CUDA(cudaStreamWaitEvent(k->stream, r->evFork, 0));
        rc = render_impl(k, cam_pos);
        if (rc != NRB_OK) return rc;
        NRB_CUDA(cudaEventRecord(k->evDone, k->stream));
        NRB_CUDA(cudaStreamWaitEvent(s, k->evDone, 0));
        launches += k->lastLaunches;
        rays += k->lastRayCount;
    }
    if (r->gather.role != 0) {
        ++r->gather.frames;
        if (r->gather.role == 2) launch_gather_signal(static_cast<unsigned*>(r->gather.remoteSync), s);
        else launch_gather_wait(r->gather.sync.as<unsigned>(), r->gather.frames * static_cast<unsigned>(r->gather.peers),
                                r->gather.sync.as<unsigned>() + 4, s);
        launches += 2;
    }
    NRB_CUDA(cudaEventRecord(r->evStop, s));
    NRB_CUDA(cudaGetLastError());
    r->accumSampleCount = r->kids[0]->accumSampleCount;
    r->lastNeural = r->kids[0]->lastNeural;
    r->lastLaunches = launches;
    r->lastRayCount = rays;
    NRB_CUDA(cudaEventRecord(r->evJoin, s));
    NRB_CUDA(cudaStreamWaitEvent(cudaStreamLegacy, r->evJoin, 0));
    return NRB_OK;
}

int nrb_read_pixels(nrb_renderer* r, void* rgba, size_t bytes) {
    if (!r || !rgba) return fail(NRB_ERR_INVALID, "null argument");
    if (bytes != r->pixelCount * 4 || !r->pixels.p) return fail(NRB_ERR_SIZE, "pixel buffer size mismatch");
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaMemcpyAsync(rgba, r->pixels.p, bytes, cudaMemcpyDeviceToHost, r->stream));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    return NRB_OK;
}

// Overlapped read-back for callers that stream frames (viewer, bench e2e): the frame is copied to a
// staging buffer on the render stream (8 MB device-to-device, microseconds) and from there to the
// host on a second stream, so the PCIe transfer of frame k runs under the kernels of frame k+1.
// The staging buffer is reused only after the previous transfer has completed (event wait).
int nrb_read_pixels_async(nrb_renderer* r, void* rgba, size_t bytes) {
    if (!r || !rgba) return fail(NRB_ERR_INVALID, "null argument");
    if (bytes != r->pixelCount * 4 || !r->pixels.p) return fail(NRB_ERR_SIZE, "pixel buffer size mismatch");
    NRB_CUDA(cudaSetDevice(r->device));
    if (!r->copyStream) {
        NRB_CUDA(cudaStreamCreateWithFlags(&r->copyStream, cudaStreamNonBlocking));
        NRB_CUDA(cudaEventCreateWithFlags(&r->evStaged, cudaEventDisableTiming));
        NRB_CUDA(cudaEventCreateWithFlags(&r->evCopied, cudaEventDisableTiming));
    }
    if (r->pixelsStaged.bytes != bytes) {
        NRB_CUDA(cudaStreamSynchronize(r->copyStream));
        NRB_CUDA(cudaStreamSynchronize(r->stream));
        NRB_CUDA(r->pixelsStaged.alloc(bytes));
        r->copyPending = false;
    }
    if (r->copyPending) NRB_CUDA(cudaStreamWaitEvent(r->stream, r->evCopied, 0));
    NRB_CUDA(cudaMemcpyAsync(r->pixelsStaged.p, r->pixels.p, bytes, cudaMemcpyDeviceToDevice, r->stream));
    NRB_CUDA(cudaEventRecord(r->evStaged, r->stream));
    NRB_CUDA(cudaStreamWaitEvent(r->copyStream, r->evStaged, 0));
    NRB_CUDA(cudaMemcpyAsync(rgba, r->pixelsStaged.p, bytes, cudaMemcpyDeviceToHost, r->copyStream));
    NRB_CUDA(cudaEventRecord(r->evCopied, r->copyStream));
    // NOT joined into the legacy default stream: nrb_render orders itself behind that stream (the caller's reads of
    // devicePixels()), and a transfer joined there would serialise the next frame behind this PCIe copy.  The copy
    // reads the staging buffer, never the frame buffer, so the next frame is free to start; nrb_sync() waits for it.
    r->copyPending = true;
    return NRB_OK;
}

// ---- frame gather over peer memory ------------------------------------------------------------------
// One process per GPU.  The root rank exports its frame buffer and a sync block as CUDA IPC handles; the
// other ranks open them and from then on their k_finalize stores the row bands they own straight into the
// root's buffer over NVLink, followed by one system-scope signal.  The root's nrb_render ends with a wait
// for the signals of all peers, so when it returns (stream-ordered) nrb_device_pixels() on the root holds
// the whole frame: the gather is fused into the last kernel of the frame, no collective call follows.
int nrb_gather_export(nrb_renderer* r, int peers, void* handle, size_t handle_bytes) {
    if (!r || !handle || peers < 0) return fail(NRB_ERR_INVALID, "bad argument");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    if (handle_bytes < NRB_GATHER_HANDLE_BYTES) return fail(NRB_ERR_SIZE, "gather handle buffer too small");
    if (r->width <= 0 || r->height <= 0) return fail(NRB_ERR_INVALID, "nrb_resize first");
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    gather_reset(r);
    const size_t framePixels = static_cast<size_t>(r->width) * r->height;
    if (framePixels != r->pixelCount) {
        NRB_CUDA(r->pixels.alloc(framePixels * sizeof(uchar4)));
        NRB_CUDA(cudaMemset(r->pixels.p, 0, r->pixels.bytes));
        r->pixelCount = framePixels;
    }
    NRB_CUDA(r->gather.sync.alloc(64));
    NRB_CUDA(cudaMemset(r->gather.sync.p, 0, 64));
    cudaIpcMemHandle_t h[2];
    NRB_CUDA(cudaIpcGetMemHandle(&h[0], r->pixels.p));
    NRB_CUDA(cudaIpcGetMemHandle(&h[1], r->gather.sync.p));
    std::memcpy(handle, h, sizeof(h));
    const uint64_t dims[2] = {static_cast<uint64_t>(r->width), static_cast<uint64_t>(r->height)};
    std::memcpy(static_cast<char*>(handle) + sizeof(h), dims, sizeof(dims));
    r->gather.role = 1;
    r->gather.peers = peers;
    return NRB_OK;
}

int nrb_gather_attach(nrb_renderer* r, const void* handle, size_t handle_bytes) {
    if (!r || !handle) return fail(NRB_ERR_INVALID, "bad argument");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    if (handle_bytes < NRB_GATHER_HANDLE_BYTES) return fail(NRB_ERR_SIZE, "gather handle buffer too small");
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    gather_reset(r);
    cudaIpcMemHandle_t h[2];
    uint64_t dims[2];
    std::memcpy(h, handle, sizeof(h));
    std::memcpy(dims, static_cast<const char*>(handle) + sizeof(h), sizeof(dims));
    if (dims[0] != static_cast<uint64_t>(r->width) || dims[1] != static_cast<uint64_t>(r->height))
        return fail(NRB_ERR_SIZE, "root frame size differs from this renderer's");
    NRB_CUDA(cudaIpcOpenMemHandle(&r->gather.remotePixels, h[0], cudaIpcMemLazyEnablePeerAccess));
    cudaError_t e = cudaIpcOpenMemHandle(&r->gather.remoteSync, h[1], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        gather_reset(r);
        return fail(NRB_ERR_CUDA, std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e));
    }
    NRB_CUDA(r->gather.sync.alloc(64));
    NRB_CUDA(cudaMemset(r->gather.sync.p, 0, 64));
    r->gather.role = 2;
    return NRB_OK;
}

int nrb_gather_detach(nrb_renderer* r) {
    if (!r) return fail(NRB_ERR_INVALID, "null renderer");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    gather_reset(r);
    return NRB_OK;
}

static int fetch_stats(nrb_renderer* r) {
    if (r->lastSliced && !r->kids.empty()) {
        for (uint64_t& v : r->lastStats) v = 0;
        for (nrb_renderer* k : r->kids) {
            int rc = fetch_stats(k);
            if (rc != NRB_OK) return rc;
            for (int i = 0; i < 8; ++i) r->lastStats[i] += k->lastStats[i];
        }
        return NRB_OK;
    }
    if (!r->st.stats) return NRB_OK;
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    NRB_CUDA(cudaMemcpy(r->lastStats, r->st.stats, 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return NRB_OK;
}

int nrb_get_timings(nrb_renderer* r, double ms[NRB_TIMING_BUCKETS], int* ray_count, uint64_t* neural_ray_calls) {
    if (!r || !ms) return fail(NRB_ERR_INVALID, "null argument");
    int rc = fetch_stats(r);
    if (rc != NRB_OK) return rc;
    for (int i = 0; i < NRB_TIMING_BUCKETS; ++i) ms[i] = 0.0;
    float t = 0.0f;
    if (r->profile || r->profPending) {
        // NRB_OPT_PROFILE (the reference's -DPROFILE_KERNELS): per-kernel events of the last frame, by bucket
        rc = drain_profiler(r);
        if (rc != NRB_OK) return rc;
        if (r->lastBucketsValid)
            for (int i = 0; i < NRB_TIMING_BUCKETS; ++i) ms[i] = r->lastBuckets[i];
    } else if (r->lastLaunches > 0 && cudaEventElapsedTime(&t, r->evStart, r->evStop) == cudaSuccess) {
        // stages are fused across the reference's buckets; the frame total is attributed to the
        // traversal bucket of the active mode (KID_primaryTrace / KID_shellTrace)
        ms[r->lastNeural ? 2 : 0] = t;
    }
    if (ray_count) *ray_count = r->lastRayCount;
    if (neural_ray_calls) *neural_ray_calls = r->lastStats[1];
    return NRB_OK;
}

int nrb_get_stage(nrb_renderer* r, int index, const char** name, double* total_ms, uint64_t* launches) {
    if (!r || index < 0) return fail(NRB_ERR_INVALID, "bad argument");
    NRB_CUDA(cudaSetDevice(r->device));
    int rc = drain_profiler(r);
    if (rc != NRB_OK) return rc;
    if (index >= static_cast<int>(r->stageOrder.size())) return 1;  // past the end
    const auto& e = *r->stages.find(r->stageOrder[index]);
    if (name) *name = e.first.c_str();
    if (total_ms) *total_ms = e.second.first;
    if (launches) *launches = e.second.second;
    return NRB_OK;
}

int nrb_reset_stages(nrb_renderer* r) {
    if (!r) return fail(NRB_ERR_INVALID, "null renderer");
    int rc = drain_profiler(r);
    r->stages.clear();
    r->stageOrder.clear();
    return rc;
}

int nrb_get_trace_stats(nrb_renderer* r, uint64_t out[3]) {
    if (!r || !out) return fail(NRB_ERR_INVALID, "null argument");
    out[0] = r->lastStats[0];  // rays of the last nrb_trace_rays_host call
    out[1] = r->lastStats[3];  // node pairs fetched
    out[2] = r->lastStats[4];  // triangle blocks fetched
    return NRB_OK;
}

int nrb_get_counters(nrb_renderer* r, uint64_t out[4]) {
    if (!r || !out) return fail(NRB_ERR_INVALID, "null argument");
    int rc = fetch_stats(r);
    if (rc != NRB_OK) return rc;
    out[0] = static_cast<uint64_t>(r->lastLaunches);
    out[1] = r->lastStats[0];
    out[2] = r->lastStats[1];
    out[3] = r->lastStats[2];
    return NRB_OK;
}

}  // extern "C"

// =================================================================================================
// Parity/debug read-back and the batch entry points
// =================================================================================================
extern "C" {

int nrb_read_buffer(nrb_renderer* r, const char* name, void* dst, size_t bytes) {
    if (!r || !name || !dst) return fail(NRB_ERR_INVALID, "null argument");
    if (r->lastSliced) return fail(NRB_ERR_STATE, "the last frame ran as concurrent slices: set NRB_OPT_CAPTURE = 1 before the frame to read its per-element buffers");
    if (!r->stateElems) return fail(NRB_ERR_STATE, "no frame has been rendered");
    NRB_CUDA(cudaSetDevice(r->device));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    const std::string n(name);
    const size_t elems = r->stateElems;
    if (n == "accum") {
        if (bytes != r->accum.bytes) return fail(NRB_ERR_SIZE, "buffer size mismatch for accum");
        NRB_CUDA(cudaMemcpy(dst, r->accum.p, bytes, cudaMemcpyDeviceToHost));
        return NRB_OK;
    }
    if (n == "tile_cost") {  // (development aid) clocks per 8x4 primary tile of the last ground-truth frame that measured them
        if (!r->tileCost.p || bytes > r->tileCost.bytes) return fail(NRB_ERR_SIZE, "buffer size mismatch for tile_cost");
        NRB_CUDA(cudaMemcpy(dst, r->tileCost.p, bytes, cudaMemcpyDeviceToHost));
        return NRB_OK;
    }
    // element buffers are repacked on the host from the 16-byte records into the reference's
    // triplet/scalar arrays (src/cuda_renderer_neural.h:160-214); debug path, not timed
    const SurfRec& rec = r->st.rec[0];
    const float4* src4 = nullptr;
    const float2* src2 = nullptr;
    int mode = 0;  // 0: xyz of src4; 1: w of src4 as float; 2: w of src4 as int; 3: (C.w, M.x, M.y)
    if (n == "hit_positions") src4 = rec.P;
    else if (n == "hit_normals") src4 = rec.N;
    else if (n == "hit_colors") src4 = rec.C;
    else if (n == "path_radiance") src4 = r->st.L;
    else if (n == "hit_distances") { src4 = rec.N; mode = 1; }
    else if (n == "hit_flags") { src4 = rec.P; mode = 2; }
    else if (n == "hit_material_params") { src4 = rec.C; src2 = rec.M; mode = 3; }
    else return fail(NRB_ERR_INVALID, "unknown buffer " + n);
    const size_t per = (mode == 1 || mode == 2) ? 1 : 3;
    if (bytes != elems * per * 4) return fail(NRB_ERR_SIZE, "buffer size mismatch for " + n);
    std::vector<float4> h4(elems);
    NRB_CUDA(cudaMemcpy(h4.data(), src4, elems * sizeof(float4), cudaMemcpyDeviceToHost));
    std::vector<float2> h2;
    if (src2) {
        h2.resize(elems);
        NRB_CUDA(cudaMemcpy(h2.data(), src2, elems * sizeof(float2), cudaMemcpyDeviceToHost));
    }
    float* out = static_cast<float*>(dst);
    for (size_t i = 0; i < elems; ++i) {
        switch (mode) {
            case 0: out[3 * i] = h4[i].x; out[3 * i + 1] = h4[i].y; out[3 * i + 2] = h4[i].z; break;
            case 1: out[i] = h4[i].w; break;
            case 2: std::memcpy(&out[i], &h4[i].w, 4); break;
            default: out[3 * i] = h4[i].w; out[3 * i + 1] = h2[i].x; out[3 * i + 2] = h2[i].y; break;
        }
    }
    return NRB_OK;
}

int nrb_trace_rays(nrb_renderer* r, int slot, int mode, const float* d_origins, const float* d_directions, size_t n,
                   nrb_hit* d_out) {
    if (!r || slot < 0 || slot >= NRB_MESH_SLOT_COUNT || mode < 0 || mode > 10) return fail(NRB_ERR_INVALID, "bad argument");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    if (n >= (1ull << 31)) return fail(NRB_ERR_INVALID, "too many rays for one call");
    NRB_CUDA(cudaSetDevice(r->device));
    if (!r->scratchC.p) {
        NRB_CUDA(r->scratchC.alloc(16 * sizeof(unsigned long long)));
    }
    NRB_CUDA(cudaMemsetAsync(r->scratchC.p, 0, r->scratchC.bytes, r->stream));
    launch_trace_rays(r->mesh[slot].view, mode, d_origins, d_directions, n, reinterpret_cast<Hit*>(d_out),
                      r->scratchC.as<unsigned long long>(), r->stream);
    NRB_CUDA(cudaGetLastError());
    return NRB_OK;
}

int nrb_trace_rays_host(nrb_renderer* r, int slot, int mode, const float* origins, const float* directions, size_t n,
                        nrb_hit* out) {
    if (!r || !origins || !directions || !out) return fail(NRB_ERR_INVALID, "null argument");
    if (n == 0) return NRB_OK;
    NRB_CUDA(cudaSetDevice(r->device));
    DevBuf o, d, h;
    NRB_CUDA(o.alloc(n * 12)); NRB_CUDA(d.alloc(n * 12)); NRB_CUDA(h.alloc(n * sizeof(nrb_hit)));
    NRB_CUDA(cudaMemcpyAsync(o.p, origins, n * 12, cudaMemcpyHostToDevice, r->stream));
    NRB_CUDA(cudaMemcpyAsync(d.p, directions, n * 12, cudaMemcpyHostToDevice, r->stream));
    int rc = nrb_trace_rays(r, slot, mode, o.as<float>(), d.as<float>(), n, h.as<nrb_hit>());
    if (rc != NRB_OK) return rc;
    NRB_CUDA(cudaMemcpyAsync(out, h.p, n * sizeof(nrb_hit), cudaMemcpyDeviceToHost, r->stream));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    NRB_CUDA(cudaMemcpy(r->lastStats, r->scratchC.p, 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
    return NRB_OK;
}

int nrb_neural_query(nrb_renderer* r, const float* d_inputs, size_t n, void* d_outputs_fp16) {
    if (!r || !d_inputs || !d_outputs_fp16) return fail(NRB_ERR_INVALID, "null argument");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    if (!r->weightsLoaded) return fail(NRB_ERR_STATE, "no weights loaded");
    if (n >= (1ull << 31)) return fail(NRB_ERR_INVALID, "too many queries for one call");
    NRB_CUDA(cudaSetDevice(r->device));
    {
        const int rcx = ensure_xt(r, n);
        if (rcx != NRB_OK) return rcx;
    }
    if (launch_neural_query(r->net, d_inputs, n, nullptr, static_cast<__half*>(d_outputs_fp16), r->stream) < 0)
        return fail(NRB_ERR_INVALID, "unsupported network shape");
    NRB_CUDA(cudaGetLastError());
    return NRB_OK;
}

int nrb_neural_query_static(nrb_renderer* r, const float* d_inputs, size_t n, void* d_outputs_fp16) {
    if (!r || !d_inputs || !d_outputs_fp16) return fail(NRB_ERR_INVALID, "null argument");
    r->touched = true;  // the next pipelined frame waits for what this call enqueues
    if (!r->weightsLoaded) return fail(NRB_ERR_STATE, "no weights loaded");
    if (n > split_static_max_rows())
        return fail(NRB_ERR_INVALID, "nrb_neural_query_static: n exceeds NRB_STATIC_QUERY_MAX (16384 queries per call)");
    NRB_CUDA(cudaSetDevice(r->device));
    const char* why = nullptr;
    const int rc = launch_split_query_static(r->net, d_inputs, n, static_cast<__half*>(d_outputs_fp16), r->stream, &why);
    if (rc < 0) return fail(rc == -2 ? NRB_ERR_CUDA : NRB_ERR_INVALID, why ? why : "static-scratch query refused");
    NRB_CUDA(cudaGetLastError());
    return NRB_OK;
}

int nrb_neural_query_host(nrb_renderer* r, const float* inputs, size_t n, void* outputs_fp16) {
    if (!r || !inputs || !outputs_fp16) return fail(NRB_ERR_INVALID, "null argument");
    if (n == 0) return NRB_OK;
    NRB_CUDA(cudaSetDevice(r->device));
    const size_t inBytes = n * r->net.inputDims * sizeof(float), outBytes = n * 16 * sizeof(__half);
    if (r->scratchA.bytes < inBytes) NRB_CUDA(r->scratchA.alloc(inBytes));
    if (r->scratchB.bytes < outBytes) NRB_CUDA(r->scratchB.alloc(outBytes));
    NRB_CUDA(cudaMemcpyAsync(r->scratchA.p, inputs, inBytes, cudaMemcpyHostToDevice, r->stream));
    int rc = nrb_neural_query(r, r->scratchA.as<float>(), n, r->scratchB.p);
    if (rc != NRB_OK) return rc;
    NRB_CUDA(cudaMemcpyAsync(outputs_fp16, r->scratchB.p, outBytes, cudaMemcpyDeviceToHost, r->stream));
    NRB_CUDA(cudaStreamSynchronize(r->stream));
    return NRB_OK;
}

}  // extern "C"
