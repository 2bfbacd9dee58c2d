// oracle/frame_oracle.cpp — TEST INFRASTRUCTURE (CPU oracle), never linked into libnrb200.so.
//
// CPU restatement of the reference's per-frame path: every __global__ kernel body of
// /root/reference/src/cuda_renderer_neural.cu:244-1607 as a plain loop over the same element
// index, and the host orchestration of RendererNeural::render / traceNeuralSegmentsForRays
// (:1810-2755) around them, on buffers in the reference's layout (AoS fp32 triplets indexed
// by sampleIdx = pixelIdx + s * pixelCount).  The per-ray arithmetic (traversal, triangle
// test, hit data, materials, RNG, ray generation, Disney BRDF) comes from backend.h: the
// reference's own headers in the oracle/_ref build, our restatement otherwise.  The neural
// query (tiny-cuda-nn has no CPU path) is restated in neural_oracle.cpp.
//
// Pinning: the reference ships no golden vectors for this path, so the oracle
// is pinned by running the reference's own headers (the _ref build) beside the restatement on
// the same inputs (tests/test_oracle_pinning.py) plus the external known answers of
// ext/bvh/test/simple_example.cpp and tiny-cuda-nn/tests/test_grid.cu.  The neural query has
// no runnable reference in this container: parity for it is UNPINNED.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load the library built from this file.

#include "backend.h"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>
#include <thread>
#include <vector>

#include "../include/nrb200.h"
#include "neural_oracle.h"

namespace {

struct Px4 {
    unsigned char x, y, z, w;
};

// ---------------------------------------------------------------------------------------------
// File-local helpers of cuda_renderer_neural.cu (anonymous namespace there, so restated here).
// ---------------------------------------------------------------------------------------------

// cuda_renderer_neural.cu:82-84
inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
// :86-88
inline float safeAcosf(float v) { return acosf(clampf(v, -1.0f, 1.0f)); }
// :90-92
inline Vec3 mul3(Vec3 a, Vec3 b) { return Vec3(a.x * b.x, a.y * b.y, a.z * b.z); }

// :94-108
inline float linearToSrgb(float v) {
    v = fmaxf(0.0f, v);
    float result;
    if (v <= 0.0031308f) {
        result = 12.92f * v;
    } else {
        result = 1.055f * powf(v, 1.0f / 2.4f) - 0.055f;
    }
    return fminf(1.0f, result);
}

// :111-113
inline Vec3 encodeSrgb(Vec3 c) { return Vec3(linearToSrgb(c.x), linearToSrgb(c.y), linearToSrgb(c.z)); }

// :115-171  (equirect lookup, manual bilinear with clamp, strength, clamp 100; else sky gradient)
inline Vec3 sampleEnvironment(EnvironmentDeviceView env, Vec3 dir) {
    if (env.pixels && env.width > 0 && env.height > 0) {
        if (env.rotation != 0.0f) {
            const float kDegToRad = 3.14159265358979323846f / 180.0f;
            float angle = env.rotation * kDegToRad;
            float cosA = cosf(angle);
            float sinA = sinf(angle);
            float newX = dir.x * cosA + dir.z * sinA;
            float newZ = -dir.x * sinA + dir.z * cosA;
            dir.x = newX;
            dir.z = newZ;
        }
        const float kInvPi = 0.3183098861837907f;
        float u = atan2f(dir.x, -dir.z) * kInvPi * 0.5f + 0.5f;
        float v = safeAcosf(dir.y) * kInvPi;
        u = clampf(u, 0.0f, 1.0f);
        v = clampf(v, 0.0f, 1.0f);
        float x = u * static_cast<float>(env.width) - 0.5f;
        float y = v * static_cast<float>(env.height) - 0.5f;
        int x0 = static_cast<int>(floorf(x));
        int y0 = static_cast<int>(floorf(y));
        int x1 = x0 + 1;
        int y1 = y0 + 1;
        auto clampi = [](int a, int n) { return a < 0 ? 0 : (a >= n ? n - 1 : a); };
        x0 = clampi(x0, env.width);
        y0 = clampi(y0, env.height);
        x1 = clampi(x1, env.width);
        y1 = clampi(y1, env.height);
        float tx = x - static_cast<float>(x0);
        float ty = y - static_cast<float>(y0);
        Vec3 c00 = env.pixels[y0 * env.width + x0];
        Vec3 c10 = env.pixels[y0 * env.width + x1];
        Vec3 c01 = env.pixels[y1 * env.width + x0];
        Vec3 c11 = env.pixels[y1 * env.width + x1];
        Vec3 c0 = c00 * (1.0f - tx) + c10 * tx;
        Vec3 c1 = c01 * (1.0f - tx) + c11 * tx;
        Vec3 envColor = (c0 * (1.0f - ty) + c1 * ty) * env.strength;
        envColor.x = fminf(envColor.x, 100.0f);
        envColor.y = fminf(envColor.y, 100.0f);
        envColor.z = fminf(envColor.z, 100.0f);
        return envColor;
    }
    float skyT = 0.5f * (dir.y + 1.0f);
    return lerp(Vec3(1.0f, 1.0f, 1.0f), Vec3(0.5f, 0.7f, 1.0f), skyT);
}

// :174-190
inline void buildTangentSpace(Vec3 normal, Vec3* tangent, Vec3* bitangent) {
    Vec3 b0(0.0f, 0.0f, 0.0f);
    if (normal.x < 0.6f && normal.x > -0.6f) {
        b0.x = 1.0f;
    } else if (normal.y < 0.6f && normal.y > -0.6f) {
        b0.y = 1.0f;
    } else if (normal.z < 0.6f && normal.z > -0.6f) {
        b0.z = 1.0f;
    } else {
        b0.x = 1.0f;
    }
    *tangent = normalize(cross(b0, normal));
    *bitangent = normalize(cross(normal, *tangent));
}

// :193-202
inline Vec3 sampleEnvironmentWithClamp(EnvironmentDeviceView env, Vec3 d, float maxRadiance) {
    Vec3 e = sampleEnvironment(env, d);
    return Vec3(fminf(e.x, maxRadiance), fminf(e.y, maxRadiance), fminf(e.z, maxRadiance));
}

struct HitData {
    bool hit;
    Vec3 position, normal, albedo, materialParams;
    float distance;
};

// Resolve (albedo, metallic/roughness/specular) at a hit: the block repeated at
// :218-229, :587-601, :1164-1178, :1276-1290.
inline void resolveAtHit(const HitInfo& h, const MeshDeviceView& mesh, const Material& global, Vec3* albedo,
                         Vec3* mparams) {
    const Material* mat = &global;
    if (h.materialId >= 0 && h.materialId < mesh.numMaterials && mesh.materials) {
        mat = &mesh.materials[h.materialId];
    }
    ResolvedMaterial r = resolveMaterial(*mat, h.uv, mesh);
    *albedo = r.base_color;
    *mparams = Vec3(r.metallic, r.roughness, r.specular);
}

inline void put3(float* buf, int base, Vec3 v) {
    buf[base + 0] = v.x;
    buf[base + 1] = v.y;
    buf[base + 2] = v.z;
}
inline Vec3 get3(const float* buf, int base) { return Vec3(buf[base + 0], buf[base + 1], buf[base + 2]); }

constexpr int kMaxSegmentIterations = 10;  // :714
constexpr float kSegmentEpsilon = 1e-8f;   // :715

// Parallel-for over [0, n) in contiguous chunks (the "grid").
struct Pool {
    int threads = 1;
    void run(int n, const std::function<void(int, int)>& fn) const {
        if (n <= 0) return;
        int t = std::max(1, std::min(threads, n));
        if (t == 1) {
            fn(0, n);
            return;
        }
        std::vector<std::thread> th;
        std::atomic<int> next{0};
        const int chunk = std::max(1, n / (t * 8));
        for (int i = 0; i < t; ++i) {
            th.emplace_back([&]() {
                for (;;) {
                    int b = next.fetch_add(chunk);
                    if (b >= n) break;
                    fn(b, std::min(n, b + chunk));
                }
            });
        }
        for (auto& x : th) x.join();
    }
};

}  // namespace

// ---------------------------------------------------------------------------------------------
// Scene storage: host copies in exactly the arrays Mesh::deviceView exposes (mesh.cu:224-261).
// ---------------------------------------------------------------------------------------------
struct OrcMesh {
    std::vector<Vec3> vertices, normals;
    std::vector<Vec2> texcoords;
    std::vector<uint3> indices;
    std::vector<BvhNode> nodes;
    std::vector<Material> materials;
    std::vector<uint32_t> materialMap;
    std::vector<int> materialIds;
    std::vector<std::vector<unsigned char>> texPixels;
    std::vector<TextureDeviceView> textures;
    int textureNearest = 0;
    int hasMeshMaterials = 0;
    Vec3 bmin, bmax;

    MeshDeviceView view() const {
        MeshDeviceView v{};
        if (indices.empty()) return v;
        v.vertices = vertices.data();
        v.normals = normals.empty() ? nullptr : normals.data();
        v.texcoords = texcoords.empty() ? nullptr : texcoords.data();
        v.indices = indices.data();
        v.numTriangles = static_cast<int>(indices.size());
        v.numVertices = static_cast<int>(vertices.size());
        v.bvhNodes = nodes.data();
        v.numBvhNodes = static_cast<int>(nodes.size());
        v.materials = materials.empty() ? nullptr : materials.data();
        v.materialMap = materialMap.empty() ? nullptr : materialMap.data();
        v.materialIds = materialIds.empty() ? nullptr : materialIds.data();
        v.numPrimitives = static_cast<int>(materialMap.size());
        v.numMaterials = static_cast<int>(materials.size());
        v.textures = textures.empty() ? nullptr : textures.data();
        v.numTextures = static_cast<int>(textures.size());
        v.textureNearest = textureNearest;
        v.hasMeshMaterials = hasMeshMaterials;
        v.hasNormals = normals.empty() ? 0 : 1;
        v.hasTexcoords = texcoords.empty() ? 0 : 1;
        return v;
    }
};

static_assert(sizeof(BvhNode) == sizeof(nrb_bvh_node), "BvhNode layout");
static_assert(sizeof(Material) == sizeof(nrb_material), "Material layout");

struct orc_renderer {
    OrcMesh mesh[4];
    std::vector<Vec3> envPixels;
    int envW = 0, envH = 0;
    float envStrength = 1.0f;
    Material globalMaterial = Material::defaultMaterial();
    OrcNetwork net;
    bool weightsLoaded = false;

    // RendererNeural state (cuda_renderer_neural.h:216-277)
    RenderBasis basis{};
    int width = 0, height = 0;
    bool lambertView = false, useNeuralQuery = false, denoise = false;
    int samplesPerPixel = 1, bounceCount = 0, classicMeshIndex = 0;
    float envmapRotation = 0.0f;
    bool useConstantNeuralColor = false;
    Vec3 constantNeuralColor{0.8f, 0.8f, 0.8f};
    bool useDirectEnvColor = false;
    Vec3 directEnvColor{0.0f, 0.0f, 0.0f};
    bool useAdditionalMesh = true;
    uint32_t accumSampleCount = 0;
    bool lastUseNeuralQuery = true, lastLambertView = false;
    int lastBounceCount = -1, lastSamplesPerPixel = -1, lastClassicMeshIndex = -1;
    bool lastUseAdditionalMesh = true;
    float lastEnvmapRotation = 0.0f;
    bool hasLastCamera = false;
    Vec3 lastCamPos{};
    RenderBasis lastBasis{};
    float lastFovY = 0.0f;
    float sceneScale = 1.0f;

    // (oracle only) restrict every per-pixel loop to rows [rowBegin,rowEnd) for bounded timing.
    int rowBegin = 0, rowEnd = 1 << 30;
    // (oracle only) of those rows, visit bands of rowBand rows, every rowStride-th band: a uniform sample of the frame
    int rowBand = 1, rowStride = 1;
    Pool pool;
    uint64_t counters[4] = {0, 0, 0, 0};  // [1] ray segments traced, [2] neural queries, [3] path vertices

    // Buffers (cuda_renderer_neural.h:150-214), element-indexed.
    std::vector<float> networkInputs;
    std::vector<uint16_t> outputs;
    std::vector<int> hitIndices;
    std::vector<float> rayDirections;
    std::vector<float> hitPositions, hitNormals, hitColors, hitMaterialParams, hitDistances;
    std::vector<int> hitFlags;
    std::vector<float> addPositions, addNormals, addColors, addMaterialParams;
    std::vector<int> addFlags;
    std::vector<float> bouncePositions, bounceNormals, bounceColors, bounceMaterialParams;
    std::vector<int> bounceHitFlags;
    std::vector<Vec3> pathThroughput, pathRadiance;
    std::vector<int> pathActive;
    std::vector<float> bounceOrigins, bounceDirections, prevBounceDirections, bouncePdfs, bounceBRDFs,
        bounceDistances;
    std::vector<int> rayActiveFlags, innerHitFlags;
    std::vector<float> accumT, currentEntryPos, outerExitT, innerEnterT, segmentExitPos;
    std::vector<Vec3> accum;
    std::vector<Px4> pixels;
    size_t bufferElements = 0;

    void ensureBuffers(size_t n, size_t npix);
    void resetAccum() {
        std::fill(accum.begin(), accum.end(), Vec3());
        accumSampleCount = 0;
    }
    EnvironmentDeviceView envView() const {
        EnvironmentDeviceView e{};
        e.pixels = envPixels.empty() ? nullptr : envPixels.data();
        e.width = envW;
        e.height = envH;
        e.rotation = envmapRotation;
        e.strength = envStrength;
        return e;
    }
    template <class F>
    void forPixels(const RenderParams& p, F body) {
        int y0 = std::max(0, rowBegin), y1 = std::min(p.height, rowEnd);
        if (rowStride <= 1) {
            pool.run(y1 - y0, [&](int a, int b) {
                for (int y = y0 + a; y < y0 + b; ++y)
                    for (int x = 0; x < p.width; ++x) body(x, y);
            });
            return;
        }
        std::vector<int> rows;
        for (int y = y0; y < y1; ++y)
            if (((y - y0) / rowBand) % rowStride == 0) rows.push_back(y);
        pool.run(static_cast<int>(rows.size()), [&](int a, int b) {
            for (int k = a; k < b; ++k)
                for (int x = 0; x < p.width; ++x) body(x, rows[k]);
        });
    }
    void traceNeuralSegmentsForRays(bool useCameraRays, const float* rayOrigins, const float* rayDirs,
                                    const int* rayActiveMask, const float* rayPdfs, size_t elementCount,
                                    const RenderParams& params, const MeshDeviceView& outerView,
                                    const MeshDeviceView& innerView, Vec3 outerMin, Vec3 outerInvExtent,
                                    float* outHitPositions, float* outHitNormals, float* outHitColors,
                                    float* outHitMaterialParams, int* outHitFlags, float* outHitDistances);
    void render(Vec3 camPos);
};

// ensureNetworkBuffers / ensureAccumBuffer (cuda_renderer_neural.cu:2826-2975)
void orc_renderer::ensureBuffers(size_t n, size_t npix) {
    if (n != bufferElements) {
        auto f3 = [&](std::vector<float>& v) { v.assign(n * 3, 0.0f); };
        auto f1 = [&](std::vector<float>& v) { v.assign(n, 0.0f); };
        auto i1 = [&](std::vector<int>& v) { v.assign(n, 0); };
        networkInputs.assign(n * net.inputDims(), 0.0f);
        outputs.assign(n * 16, 0);
        i1(hitIndices);
        f3(rayDirections);
        f3(hitPositions); f3(hitNormals); f3(hitColors); f3(hitMaterialParams); f1(hitDistances); i1(hitFlags);
        f3(addPositions); f3(addNormals); f3(addColors); f3(addMaterialParams); i1(addFlags);
        f3(bouncePositions); f3(bounceNormals); f3(bounceColors); f3(bounceMaterialParams); i1(bounceHitFlags);
        pathThroughput.assign(n, Vec3()); pathRadiance.assign(n, Vec3()); i1(pathActive);
        f3(bounceOrigins); f3(bounceDirections); f3(prevBounceDirections); f1(bouncePdfs); f3(bounceBRDFs);
        f1(bounceDistances);
        i1(rayActiveFlags); i1(innerHitFlags);
        f1(accumT); f3(currentEntryPos); f1(outerExitT); f1(innerEnterT); f3(segmentExitPos);
        bufferElements = n;
    }
    if (npix != accum.size()) {
        accum.assign(npix, Vec3());
        accumSampleCount = 0;
    }
    if (npix != pixels.size()) pixels.assign(npix, Px4{0, 0, 0, 0});
}

// ---------------------------------------------------------------------------------------------
// traceNeuralSegmentsForRays (cuda_renderer_neural.cu:1810-2083) with the kernel bodies
// traceOuterShellEntryKernel :720-767, traceOuterShellEntryFromRaysKernel :773-852,
// compactInputsKernel :1529-1541, traceSegmentExitsKernel :858-930,
// buildSegmentNeuralInputsKernel :937-1001, applySegmentNeuralOutputKernel :1007-1123,
// prepareNextIterationKernel :1456-1524 inlined as loops.
// ---------------------------------------------------------------------------------------------
void orc_renderer::traceNeuralSegmentsForRays(bool useCameraRays, const float* rayOrigins, const float* rayDirs,
                                              const int* rayActiveMask, const float* rayPdfs, size_t elementCount,
                                              const RenderParams& params, const MeshDeviceView& outerView,
                                              const MeshDeviceView& innerView, Vec3 outerMin, Vec3 outerInvExtent,
                                              float* outHitPositions, float* outHitNormals, float* outHitColors,
                                              float* outHitMaterialParams, int* outHitFlags,
                                              float* outHitDistances) {
    if (elementCount == 0) return;
    const int pointCount = net.cfg.n_points;
    const int inputDims = net.inputDims();
    const Material material = params.material;
    std::atomic<uint64_t> nTraces{0};

    // the six cudaMemsets (:1842-1851)
    std::fill(outHitFlags, outHitFlags + elementCount, 0);
    std::fill(outHitPositions, outHitPositions + elementCount * 3, 0.0f);
    std::fill(outHitNormals, outHitNormals + elementCount * 3, 0.0f);
    std::fill(outHitColors, outHitColors + elementCount * 3, 0.0f);
    if (outHitMaterialParams) std::fill(outHitMaterialParams, outHitMaterialParams + elementCount * 3, 0.0f);
    if (outHitDistances) std::fill(outHitDistances, outHitDistances + elementCount, 0.0f);
    // Rows outside [rowBegin,rowEnd) are never visited by the bounded-sample loops.
    if (rowBegin > 0 || rowEnd < params.height || rowStride > 1) std::fill(rayActiveFlags.begin(), rayActiveFlags.end(), 0);

    float* entryPositions = currentEntryPos.data();
    float* entryT = outerExitT.data();  // the reference passes outerExitT_ as entryT (:1881)
    int* activeFlags = rayActiveFlags.data();
    const float* segDirs = nullptr;

    if (useCameraRays) {
        forPixels(params, [&](int x, int y) {  // traceOuterShellEntryKernel
            int pixelIdx = y * params.width + x;
            uint64_t local = 0;
            for (int s = 0; s < params.samplesPerPixel; ++s) {
                int sampleIdx = pixelIdx + s * params.pixelCount;
                int base = sampleIdx * 3;
                uint32_t rng = initRng(pixelIdx, params.sampleOffset, s);
                Ray ray = generatePrimaryRay(x, y, params, rng);
                put3(rayDirections.data(), base, ray.direction);
                HitInfo outerHit;
                bool hitOuter =
                    traceMeshWithMode(ray, outerView, &outerHit, TraceMode::FORWARD_ONLY, false, params.material);
                ++local;
                if (hitOuter) {
                    put3(entryPositions, base, ray.at(outerHit.t));
                    entryT[sampleIdx] = outerHit.t;
                    activeFlags[sampleIdx] = 1;
                    accumT[sampleIdx] = outerHit.t;
                } else {
                    put3(entryPositions, base, Vec3());
                    entryT[sampleIdx] = 0.0f;
                    activeFlags[sampleIdx] = 0;
                    accumT[sampleIdx] = 0.0f;
                }
            }
            nTraces += local;
        });
        segDirs = rayDirections.data();
    } else {
        if (!rayOrigins || !rayDirs) return;
        forPixels(params, [&](int x, int y) {  // traceOuterShellEntryFromRaysKernel
            int pixelIdx = y * params.width + x;
            uint64_t local = 0;
            for (int s = 0; s < params.samplesPerPixel; ++s) {
                int sampleIdx = pixelIdx + s * params.pixelCount;
                int base = sampleIdx * 3;
                auto inactive = [&]() {
                    put3(entryPositions, base, Vec3());
                    entryT[sampleIdx] = 0.0f;
                    activeFlags[sampleIdx] = 0;
                    accumT[sampleIdx] = 0.0f;
                };
                if (rayActiveMask && !rayActiveMask[sampleIdx]) { inactive(); continue; }
                if (rayPdfs && rayPdfs[sampleIdx] <= 0.0f) { inactive(); continue; }
                Vec3 origin = get3(rayOrigins, base);
                Vec3 dir = get3(rayDirs, base);
                Ray ray(origin, dir);
                HitInfo outerHit;
                bool hitOuter =
                    traceMeshWithMode(ray, outerView, &outerHit, TraceMode::FORWARD_ONLY, false, params.material);
                ++local;
                float baseOffset = 0.0f;
                Vec3 entryOrigin = origin;
                if (!hitOuter) {
                    HitInfo exitHit;
                    bool hitExit = traceMeshWithMode(ray, outerView, &exitHit, TraceMode::BACKWARD_ONLY, false,
                                                     params.material);
                    ++local;
                    if (hitExit) {
                        baseOffset = exitHit.t + kSegmentEpsilon;
                        entryOrigin = origin + dir * baseOffset;
                        Ray shifted(entryOrigin, dir);
                        hitOuter = traceMeshWithMode(shifted, outerView, &outerHit, TraceMode::FORWARD_ONLY, false,
                                                     params.material);
                        ++local;
                    }
                }
                if (hitOuter) {
                    Vec3 entryPos = entryOrigin + dir * outerHit.t;
                    float totalEntryT = baseOffset + outerHit.t;
                    put3(entryPositions, base, entryPos);
                    entryT[sampleIdx] = totalEntryT;
                    activeFlags[sampleIdx] = 1;
                    accumT[sampleIdx] = totalEntryT;
                } else {
                    inactive();
                }
            }
            nTraces += local;
        });
        segDirs = rayDirs;
    }

    // compactInputsKernel (order is arbitrary in the reference; ascending here)
    auto compact = [&]() {
        int c = 0;
        for (size_t i = 0; i < elementCount; ++i)
            if (activeFlags[i]) hitIndices[c++] = static_cast<int>(i);
        return c;
    };
    int activeCount = compact();

    for (int iter = 0; iter < kMaxSegmentIterations && activeCount > 0; ++iter) {
        size_t padded = ((static_cast<size_t>(activeCount) + 255) / 256) * 256;

        pool.run(activeCount, [&](int a, int b) {  // traceSegmentExitsKernel
            for (int idx = a; idx < b; ++idx) {
                int sampleIdx = hitIndices[idx];
                int base = sampleIdx * 3;
                Vec3 entryPos = get3(entryPositions, base);
                Vec3 dir = get3(segDirs, base);
                Vec3 shiftedEntry = entryPos + dir * kSegmentEpsilon;
                Ray ray(shiftedEntry, dir);
                HitInfo outerExit;
                bool hitOuterExit =
                    traceMeshWithMode(ray, outerView, &outerExit, TraceMode::BACKWARD_ONLY, false, material);
                float exitT = hitOuterExit ? outerExit.t : kSegmentEpsilon;
                outerExitT[sampleIdx] = exitT;
                HitInfo innerHit;
                bool hitInner = traceMeshWithMode(ray, innerView, &innerHit, TraceMode::ANY, false, material);
                float innerT;
                if (hitInner) {
                    innerT = innerHit.t;
                    innerHitFlags[sampleIdx] = 1;
                } else {
                    innerT = 1e30f;
                    innerHitFlags[sampleIdx] = 0;
                }
                innerEnterT[sampleIdx] = innerT;
                bool innerBeforeOuter = hitInner && (innerT < exitT);
                Vec3 exitPos = innerBeforeOuter ? shiftedEntry + dir * innerT : shiftedEntry + dir * exitT;
                put3(segmentExitPos.data(), base, exitPos);
            }
        });
        nTraces += 2ull * activeCount;

        pool.run(activeCount, [&](int a, int b) {  // buildSegmentNeuralInputsKernel
            for (int idx = a; idx < b; ++idx) {
                int sampleIdx = hitIndices[idx];
                int srcBase = sampleIdx * 3;
                int dstBase = idx * inputDims;
                Vec3 entryPos = get3(entryPositions, srcBase);
                Vec3 dir = get3(segDirs, srcBase);
                Vec3 shiftedEntry = entryPos + dir * kSegmentEpsilon;
                Vec3 exitPos = get3(segmentExitPos.data(), srcBase);
                auto norm = [&](Vec3 p) {
                    return Vec3((p.x - outerMin.x) * outerInvExtent.x, (p.y - outerMin.y) * outerInvExtent.y,
                                (p.z - outerMin.z) * outerInvExtent.z);
                };
                put3(networkInputs.data(), dstBase + 0, norm(shiftedEntry));
                put3(networkInputs.data(), dstBase + 3, norm(exitPos));
                if (pointCount == 3) {
                    Vec3 midPos = (shiftedEntry + exitPos) * 0.5f;
                    put3(networkInputs.data(), dstBase + 6, norm(midPos));
                }
                int dirOffset = pointCount * 3;
                networkInputs[dstBase + dirOffset + 0] = (dir.x + 1.0f) * 0.5f;
                networkInputs[dstBase + dirOffset + 1] = (dir.y + 1.0f) * 0.5f;
                networkInputs[dstBase + dirOffset + 2] = (dir.z + 1.0f) * 0.5f;
            }
        });
        if (padded > static_cast<size_t>(activeCount)) {  // tail memset (:2017-2022)
            std::fill(networkInputs.begin() + static_cast<size_t>(activeCount) * inputDims,
                      networkInputs.begin() + padded * inputDims, 0.0f);
        }

        // network_->inference_mixed_precision (:2031)
        counters[2] += activeCount;
        pool.run(activeCount, [&](int a, int b) {
            net.forward(networkInputs.data() + static_cast<size_t>(a) * inputDims, b - a,
                        outputs.data() + static_cast<size_t>(a) * 16);
        });

        const int outputStride = 16;
        pool.run(activeCount, [&](int a, int b) {  // applySegmentNeuralOutputKernel
            for (int idx = a; idx < b; ++idx) {
                int sampleIdx = hitIndices[idx];
                int base = sampleIdx * 3;
                int outBase = idx * outputStride;
                float o[8];
                for (int k = 0; k < 8; ++k) o[k] = orc_half_to_float(outputs[outBase + k]);
                float hasIntersection = o[0], distance = o[1];
                Vec3 entryPos = get3(entryPositions, base);
                Vec3 dir = get3(segDirs, base);
                bool neuralHit = (hasIntersection >= 0.0f);
                bool innerHitBeforeExit =
                    (innerHitFlags[sampleIdx] != 0) && (innerEnterT[sampleIdx] < outerExitT[sampleIdx]);
                if (neuralHit || innerHitBeforeExit) {
                    Vec3 shiftedEntry = entryPos + dir * kSegmentEpsilon;
                    float clampedDist = distance * 1.0f;
                    if (innerEnterT[sampleIdx] > 0.0f) {
                        clampedDist = fminf(clampedDist, innerEnterT[sampleIdx] - kSegmentEpsilon);
                    }
                    put3(outHitPositions, base, shiftedEntry + dir * clampedDist);
                    Vec3 normal(o[2], o[3], o[4]);
                    float nlen = length(normal);
                    normal = nlen > 1e-6f ? normal / nlen : Vec3(0.0f, 1.0f, 0.0f);
                    put3(outHitNormals, base, normal);
                    Vec3 c;
                    if (params.useConstantNeuralColor) {
                        c = params.constantNeuralColor;
                    } else if (outputStride >= 8) {
                        c = Vec3(saturate(o[5]), saturate(o[6]), saturate(o[7]));
                    } else {
                        c = material.base_color.value;
                    }
                    put3(outHitColors, base, c);
                    if (outHitMaterialParams) {
                        put3(outHitMaterialParams, base,
                             Vec3(material.metallic.value, material.roughness.value, material.specular.value));
                    }
                    outHitFlags[sampleIdx] = 1;
                    activeFlags[sampleIdx] = 0;
                    if (outHitDistances) outHitDistances[sampleIdx] = distance;
                }
            }
        });

        std::atomic<uint64_t> reTraces{0};
        pool.run(activeCount, [&](int a, int b) {  // prepareNextIterationKernel
            uint64_t local = 0;
            for (int idx = a; idx < b; ++idx) {
                int sampleIdx = hitIndices[idx];
                if (activeFlags[sampleIdx] == 0) continue;
                int base = sampleIdx * 3;
                Vec3 exitPos = get3(segmentExitPos.data(), base);
                Vec3 dir = get3(segDirs, base);
                Vec3 shiftedExit = exitPos + dir * kSegmentEpsilon;
                Ray ray(shiftedExit, dir);
                HitInfo reentry;
                bool hitReentry =
                    traceMeshWithMode(ray, outerView, &reentry, TraceMode::FORWARD_ONLY, false, material);
                ++local;
                bool canContinue = hitReentry || (innerHitFlags[sampleIdx] != 0);
                if (canContinue) {
                    float reentryDist = hitReentry ? reentry.t : 0.0f;
                    put3(entryPositions, base, shiftedExit + dir * reentryDist);
                    accumT[sampleIdx] = accumT[sampleIdx] + outerExitT[sampleIdx] + reentryDist + 2.0f * kSegmentEpsilon;
                    innerEnterT[sampleIdx] = reentryDist;  // newEntryT aliases innerEnterT_ (:2069)
                    activeFlags[sampleIdx] = 1;
                } else {
                    activeFlags[sampleIdx] = 0;
                }
            }
            reTraces += local;
        });
        nTraces += reTraces.load();
        activeCount = compact();
    }
    counters[1] += nTraces.load();
}

// ---------------------------------------------------------------------------------------------
// RendererNeural::render (cuda_renderer_neural.cu:2085-2755), OptiX branches dropped.
// ---------------------------------------------------------------------------------------------
void orc_renderer::render(Vec3 camPos) {
    if (width <= 0 || height <= 0) return;
    int maxBounces = std::max(0, bounceCount);
    for (auto& c : counters) c = 0;

    OrcMesh& originalMesh = mesh[NRB_MESH_ORIGINAL];
    OrcMesh& outerShell = mesh[NRB_MESH_OUTER_SHELL];
    OrcMesh& innerShell = mesh[NRB_MESH_INNER_SHELL];
    OrcMesh* classicMesh = &originalMesh;  // :2119-2125
    if (classicMeshIndex == 1 && !innerShell.indices.empty()) classicMesh = &innerShell;
    else if (classicMeshIndex == 2 && !outerShell.indices.empty()) classicMesh = &outerShell;

    MeshDeviceView classicView = classicMesh->view();
    MeshDeviceView outerView = outerShell.view();
    MeshDeviceView innerView = innerShell.view();
    EnvironmentDeviceView env = envView();

    size_t pixelCount = static_cast<size_t>(width) * height;
    int spp = samplesPerPixel <= 0 ? 1 : samplesPerPixel;
    size_t elementCount = pixelCount * spp;
    ensureBuffers(((elementCount + 255) / 256) * 256, pixelCount);

    Vec3 outerMin = outerShell.bmin, outerMax = outerShell.bmax;  // :2171-2178
    Vec3 outerExtent = outerMax - outerMin;
    Vec3 outerInvExtent(outerExtent.x != 0.0f ? 1.0f / outerExtent.x : 0.0f,
                        outerExtent.y != 0.0f ? 1.0f / outerExtent.y : 0.0f,
                        outerExtent.z != 0.0f ? 1.0f / outerExtent.z : 0.0f);
    Vec3 se = classicMesh->bmax - classicMesh->bmin;  // :2180-2189
    sceneScale = sqrtf(se.x * se.x + se.y * se.y + se.z * se.z);
    if (sceneScale < 1e-6f) sceneScale = 1.0f;

    bool cameraMoved = !hasLastCamera;  // :2190-2217
    const float kEps = 1e-4f;
    if (!cameraMoved) {
        auto d3 = [&](Vec3 a, Vec3 b) {
            return fabsf(a.x - b.x) > kEps || fabsf(a.y - b.y) > kEps || fabsf(a.z - b.z) > kEps;
        };
        if (d3(camPos, lastCamPos) || d3(basis.forward, lastBasis.forward) || d3(basis.right, lastBasis.right) ||
            d3(basis.up, lastBasis.up) || fabsf(basis.fovY - lastFovY) > kEps)
            cameraMoved = true;
    }
    if (cameraMoved || useNeuralQuery != lastUseNeuralQuery || lambertView != lastLambertView ||
        maxBounces != lastBounceCount || spp != lastSamplesPerPixel || classicMeshIndex != lastClassicMeshIndex ||
        envmapRotation != lastEnvmapRotation || useAdditionalMesh != lastUseAdditionalMesh) {
        resetAccum();
    }
    lastUseNeuralQuery = useNeuralQuery; lastLambertView = lambertView; lastBounceCount = maxBounces;
    lastSamplesPerPixel = spp; lastClassicMeshIndex = classicMeshIndex; lastEnvmapRotation = envmapRotation;
    lastUseAdditionalMesh = useAdditionalMesh; lastCamPos = camPos; lastBasis = basis; lastFovY = basis.fovY;
    hasLastCamera = true;

    RenderParams params;  // :2232-2253
    params.camPos = camPos;
    params.camForward = basis.forward;
    params.camRight = basis.right;
    params.camUp = basis.up;
    params.lightDir = normalize(Vec3(1.0f, 1.5f, -1.0f));
    params.outerShellMin = outerMin;
    params.outerShellInvExtent = outerInvExtent;
    params.material = globalMaterial;
    params.useConstantNeuralColor = useConstantNeuralColor;
    params.constantNeuralColor = constantNeuralColor;
    params.useDirectEnvColor = useDirectEnvColor;
    params.directEnvColor = directEnvColor;
    params.fovY = basis.fovY;
    params.maxRadiance = 100.0f;
    params.sceneScale = sceneScale;
    params.maxBounces = maxBounces;
    params.width = width;
    params.height = height;
    params.pixelCount = static_cast<int>(pixelCount);
    params.samplesPerPixel = spp;
    params.sampleOffset = accumSampleCount;

    std::atomic<uint64_t> nTraces{0}, nVerts{0};

    // --- shared wavefront kernels as lambdas --------------------------------------------------
    auto initializePathState = [&]() {  // initializePathStateKernel :303-354
        forPixels(params, [&](int x, int y) {
            int pixelIdx = y * params.width + x;
            for (int s = 0; s < params.samplesPerPixel; ++s) {
                int sampleIdx = pixelIdx + s * params.pixelCount;
                uint32_t rng = initRng(pixelIdx, params.sampleOffset, s);
                Ray primaryRay = generatePrimaryRay(x, y, params, rng);
                Vec3 sampleRadiance(0.0f, 0.0f, 0.0f), sampleThroughput(1.0f, 1.0f, 1.0f);
                int isActive = 0;
                if (hitFlags[sampleIdx]) {
                    isActive = 1;  // (normal normalise/flip at :335-346 has no effect on outputs)
                } else {
                    sampleRadiance = params.useDirectEnvColor
                                         ? params.directEnvColor
                                         : sampleEnvironmentWithClamp(env, primaryRay.direction, params.maxRadiance);
                }
                pathThroughput[sampleIdx] = sampleThroughput;
                pathRadiance[sampleIdx] = sampleRadiance;
                pathActive[sampleIdx] = isActive;
            }
        });
    };

    auto sampleBounceDirections = [&](const float* hPos, const float* hNrm, const float* hCol, const float* hMat,
                                      const int* hFlags, const float* hDist, float* bDist,
                                      const float* incomingDirections) {  // :357-499
        forPixels(params, [&](int x, int y) {
            int pixelIdx = y * params.width + x;
            uint64_t local = 0;
            for (int s = 0; s < params.samplesPerPixel; ++s) {
                int sampleIdx = pixelIdx + s * params.pixelCount;
                int base = sampleIdx * 3;
                if (!hFlags[sampleIdx] || !pathActive[sampleIdx]) {
                    put3(bounceOrigins.data(), base, Vec3());
                    put3(bounceDirections.data(), base, Vec3());
                    bouncePdfs[sampleIdx] = 0.0f;
                    put3(bounceBRDFs.data(), base, Vec3());
                    if (bDist) bDist[sampleIdx] = 0.0f;
                    continue;
                }
                ++local;
                uint32_t rng = initRng(pixelIdx, params.sampleOffset, s);
                Vec3 hitPos = get3(hPos, base), normal = get3(hNrm, base), albedo = get3(hCol, base);
                float nlen = length(normal);
                normal = nlen > 0.0f ? normal / nlen : Vec3(0.0f, 1.0f, 0.0f);
                Vec3 incomingDir;
                if (incomingDirections) {
                    incomingDir = get3(incomingDirections, base);
                } else {
                    Ray primaryRay = generatePrimaryRay(x, y, params, rng);
                    incomingDir = primaryRay.direction;
                }
                if (dot(normal, incomingDir) > 0.0f) normal = normal * -1.0f;
                Vec3 wo = incomingDir * -1.0f;
                Vec3 tangent, bitangent;
                buildTangentSpace(normal, &tangent, &bitangent);
                Vec3 mp(params.material.metallic.value, params.material.roughness.value,
                        params.material.specular.value);
                if (hMat) mp = get3(hMat, base);
                Material surfaceMat = params.material;
                surfaceMat.base_color = MaterialParamVec3::constant(albedo);
                surfaceMat.metallic = MaterialParam::constant(mp.x);
                surfaceMat.roughness = MaterialParam::constant(mp.y);
                surfaceMat.specular = MaterialParam::constant(mp.z);
                float u1 = rand01(rng), u2 = rand01(rng), u3 = rand01(rng);
                float pdf;
                Vec3 wi = disney_sample_3component(surfaceMat, normal, wo, tangent, bitangent, u1, u2, u3, &pdf);
                if (pdf <= 0.0f) {
                    pathActive[sampleIdx] = 0;
                    bouncePdfs[sampleIdx] = 0.0f;
                    continue;
                }
                Vec3 f = disney_eval(surfaceMat, normal, wo, wi, tangent, bitangent);
                float cos_theta = fabsf(dot(normal, wi));
                Vec3 w = f * (cos_theta / pdf);
                if (std::isnan(w.x) || std::isnan(w.y) || std::isnan(w.z) || std::isinf(w.x) || std::isinf(w.y) ||
                    std::isinf(w.z)) {
                    pathActive[sampleIdx] = 0;
                    bouncePdfs[sampleIdx] = 0.0f;
                    continue;
                }
                float rayOffset = params.sceneScale * 1e-6f;
                put3(bounceOrigins.data(), base, hitPos + normal * rayOffset);
                put3(bounceDirections.data(), base, wi);
                bouncePdfs[sampleIdx] = pdf;
                put3(bounceBRDFs.data(), base, w);
                if (bDist && hDist) bDist[sampleIdx] = hDist[sampleIdx];
            }
            nVerts += local;
        });
    };

    auto integrateBounce = [&](int bounceIndex) {  // integrateBounceKernel :621-675
        forPixels(params, [&](int x, int y) {
            int pixelIdx = y * params.width + x;
            for (int s = 0; s < params.samplesPerPixel; ++s) {
                int sampleIdx = pixelIdx + s * params.pixelCount;
                if (!pathActive[sampleIdx]) continue;
                int base = sampleIdx * 3;
                Vec3 newT = mul3(pathThroughput[sampleIdx], get3(bounceBRDFs.data(), base));
                pathThroughput[sampleIdx] = newT;
                if (!bounceHitFlags[sampleIdx]) {
                    Vec3 envLight =
                        sampleEnvironmentWithClamp(env, get3(bounceDirections.data(), base), params.maxRadiance);
                    pathRadiance[sampleIdx] = pathRadiance[sampleIdx] + mul3(newT, envLight);
                    pathActive[sampleIdx] = 0;
                    continue;
                }
                if (bounceIndex >= params.maxBounces) {
                    pathActive[sampleIdx] = 0;
                    continue;
                }
                Vec3 tp = pathThroughput[sampleIdx];
                float survivalProb = fminf(1.0f, fmaxf(tp.x, fmaxf(tp.y, tp.z)));
                uint32_t rng = initRng(pixelIdx, params.sampleOffset + bounceIndex, s);
                if (rand01(rng) >= survivalProb) {
                    pathActive[sampleIdx] = 0;
                    continue;
                }
                pathThroughput[sampleIdx] = tp * (1.0f / survivalProb);
            }
        });
    };

    auto finalize = [&]() {  // finalizePathTracingKernel :678-709
        forPixels(params, [&](int x, int y) {
            int pixelIdx = y * params.width + x;
            Vec3 sum(0.0f, 0.0f, 0.0f);
            for (int s = 0; s < params.samplesPerPixel; ++s) sum += pathRadiance[pixelIdx + s * params.pixelCount];
            Vec3 newSum = accum[pixelIdx] + sum;
            accum[pixelIdx] = newSum;
            float invSamples = 1.0f / static_cast<float>(params.sampleOffset + params.samplesPerPixel);
            Vec3 color = encodeSrgb(newSum * invSamples);
            pixels[pixelIdx] = Px4{static_cast<unsigned char>(color.x * 255.0f),
                                   static_cast<unsigned char>(color.y * 255.0f),
                                   static_cast<unsigned char>(color.z * 255.0f), 255};
        });
    };

    // bilateralDenoiseKernel, src/denoiser.cuh:47-113 (shipped by the reference, never launched there):
    // guides = sample-0 primary-hit normals / albedo, accumCount = samples accumulated so far.
    auto denoisePass = [&]() {
        const int kRadius = 5;                                     // denoiser.cuh:31
        const float kSigmaSpatial = 3.0f, kNormalAlpha = 64.0f, kSigmaAlbedo = 0.1f;  // :32-34
        const int width = params.width, height = params.height;
        const int accumCount = static_cast<int>(params.sampleOffset) + params.samplesPerPixel;
        const float* guideNormals = hitNormals.data();
        const float* guideAlbedo = hitColors.data();
        forPixels(params, [&](int cx, int cy) {
            int centerIdx = cy * width + cx;
            float invAccum = 1.0f / fmaxf(1.0f, static_cast<float>(accumCount));
            Vec3 centerColor = accum[centerIdx] * invAccum;
            float cnx = guideNormals[centerIdx * 3 + 0], cny = guideNormals[centerIdx * 3 + 1], cnz = guideNormals[centerIdx * 3 + 2];
            float cax = guideAlbedo[centerIdx * 3 + 0], cay = guideAlbedo[centerIdx * 3 + 1], caz = guideAlbedo[centerIdx * 3 + 2];
            const float invSigmaSpatial2 = 1.0f / (2.0f * kSigmaSpatial * kSigmaSpatial);
            const float invSigmaAlbedo2 = 1.0f / (2.0f * kSigmaAlbedo * kSigmaAlbedo);
            Vec3 weightedSum(0.0f, 0.0f, 0.0f);
            float totalWeight = 0.0f;
            for (int dy = -kRadius; dy <= kRadius; ++dy) {
                int ny = cy + dy;
                if (ny < 0 || ny >= height) continue;
                for (int dx = -kRadius; dx <= kRadius; ++dx) {
                    int nx = cx + dx;
                    if (nx < 0 || nx >= width) continue;
                    int nIdx = ny * width + nx;
                    Vec3 neighborColor = accum[nIdx] * invAccum;
                    float dist2 = static_cast<float>(dx * dx + dy * dy);
                    float wSpatial = expf(-dist2 * invSigmaSpatial2);
                    float nnx = guideNormals[nIdx * 3 + 0], nny = guideNormals[nIdx * 3 + 1], nnz = guideNormals[nIdx * 3 + 2];
                    float normalDot = fmaxf(0.0f, fminf(1.0f, cnx * nnx + cny * nny + cnz * nnz));
                    float wNormal = powf(normalDot, kNormalAlpha);
                    float dax = guideAlbedo[nIdx * 3 + 0] - cax, day = guideAlbedo[nIdx * 3 + 1] - cay, daz = guideAlbedo[nIdx * 3 + 2] - caz;
                    float wAlbedo = expf(-(dax * dax + day * day + daz * daz) * invSigmaAlbedo2);
                    float w = wSpatial * wNormal * wAlbedo;
                    weightedSum += neighborColor * w;
                    totalWeight += w;
                }
            }
            Vec3 color = (totalWeight > 0.0f) ? (weightedSum * (1.0f / totalWeight)) : centerColor;
            color = encodeSrgb(color);
            pixels[centerIdx] = Px4{static_cast<unsigned char>(color.x * 255.0f), static_cast<unsigned char>(color.y * 255.0f),
                                    static_cast<unsigned char>(color.z * 255.0f), 255};
        });
    };

    auto lambert = [&]() {  // lambertKernel :1547-1607
        forPixels(params, [&](int x, int y) {
            int pixelIdx = y * params.width + x;
            Vec3 sum(0.0f, 0.0f, 0.0f);
            for (int s = 0; s < params.samplesPerPixel; ++s) {
                int sampleIdx = pixelIdx + s * params.pixelCount;
                uint32_t rng = initRng(pixelIdx, params.sampleOffset, s);
                Ray primaryRay = generatePrimaryRay(x, y, params, rng);
                Vec3 color(0.0f, 0.0f, 0.0f);
                if (hitFlags[sampleIdx]) {
                    Vec3 normal = get3(hitNormals.data(), sampleIdx * 3);
                    Vec3 baseColor = get3(hitColors.data(), sampleIdx * 3);
                    float nlen = length(normal);
                    normal = nlen > 0.0f ? normal / nlen : Vec3(0.0f, 1.0f, 0.0f);
                    if (dot(normal, primaryRay.direction) > 0.0f) normal = normal * -1.0f;
                    float ndotl = fmaxf(0.0f, dot(normal, -primaryRay.direction));
                    color = baseColor * ndotl;
                } else {
                    color = sampleEnvironment(env, primaryRay.direction);
                }
                sum += color;
            }
            Vec3 color = encodeSrgb(sum * (1.0f / static_cast<float>(params.samplesPerPixel)));
            pixels[pixelIdx] = Px4{static_cast<unsigned char>(color.x * 255.0f),
                                   static_cast<unsigned char>(color.y * 255.0f),
                                   static_cast<unsigned char>(color.z * 255.0f), 255};
        });
    };

    // Trace `mesh` for one element and write the reference's five hit buffers
    // (the shared tail of :244-296, :556-618, :1128-1188, :1235-1302).
    auto writeMeshHit = [&](const Ray& ray, const MeshDeviceView& m, bool cull, int sampleIdx, float* oPos,
                            float* oNrm, float* oCol, float* oMat, int* oFlags, bool zeroOnMiss) {
        int base = sampleIdx * 3;
        HitInfo h;
        bool hit = traceMesh(ray, m, &h, cull, params.material);
        if (hit) {
            put3(oPos, base, h.position);
            put3(oNrm, base, h.shadingNormal);
            Vec3 albedo, mp;
            resolveAtHit(h, m, params.material, &albedo, &mp);
            put3(oCol, base, albedo);
            put3(oMat, base, mp);
            oFlags[sampleIdx] = 1;
        } else {
            if (zeroOnMiss) {
                put3(oPos, base, Vec3());
                put3(oNrm, base, Vec3());
                put3(oCol, base, Vec3());
            }
            put3(oMat, base, Vec3(params.material.metallic.value, params.material.roughness.value,
                                  params.material.specular.value));
            oFlags[sampleIdx] = 0;
        }
    };

    // selectClosestPrimaryHitKernel :1307-1378 / selectClosestHitKernel :1383-1450
    auto selectClosest = [&](float* sPos, float* sNrm, float* sCol, float* sMat, int* sFlags,
                             const float* rayOrigins /* null = camera */) {
        forPixels(params, [&](int x, int y) {
            int pixelIdx = y * params.width + x;
            for (int s = 0; s < params.samplesPerPixel; ++s) {
                int sampleIdx = pixelIdx + s * params.pixelCount;
                int base = sampleIdx * 3;
                bool shellHit = sFlags[sampleIdx] != 0;
                bool addHit = addFlags[sampleIdx] != 0;
                auto take = [&]() {
                    put3(sPos, base, get3(addPositions.data(), base));
                    put3(sNrm, base, get3(addNormals.data(), base));
                    put3(sCol, base, get3(addColors.data(), base));
                    put3(sMat, base, get3(addMaterialParams.data(), base));
                };
                if (shellHit && addHit) {
                    Vec3 o = rayOrigins ? get3(rayOrigins, base) : params.camPos;
                    float shellDist = length(get3(sPos, base) - o);
                    float addDist = length(get3(addPositions.data(), base) - o);
                    if (addDist < shellDist) take();
                } else if (addHit) {
                    take();
                    sFlags[sampleIdx] = 1;
                }
            }
        });
    };

    bool neuralReady = useNeuralQuery && weightsLoaded && !outerShell.indices.empty();
    if (neuralReady) {
        traceNeuralSegmentsForRays(true, nullptr, nullptr, nullptr, nullptr, elementCount, params, outerView,
                                   innerView, outerMin, outerInvExtent, hitPositions.data(), hitNormals.data(),
                                   hitColors.data(), hitMaterialParams.data(), hitFlags.data(), hitDistances.data());
        MeshDeviceView additionalView = useAdditionalMesh ? mesh[NRB_MESH_ADDITIONAL].view() : MeshDeviceView{};

        forPixels(params, [&](int x, int y) {  // traceAdditionalMeshPrimaryRaysKernel :1128-1188
            int pixelIdx = y * params.width + x;
            uint64_t local = 0;
            for (int s = 0; s < params.samplesPerPixel; ++s) {
                int sampleIdx = pixelIdx + s * params.pixelCount;
                if (additionalView.numTriangles == 0) {
                    addFlags[sampleIdx] = 0;
                    continue;
                }
                uint32_t rng = initRng(pixelIdx, params.sampleOffset, s);
                Ray ray = generatePrimaryRay(x, y, params, rng);
                writeMeshHit(ray, additionalView, true, sampleIdx, addPositions.data(), addNormals.data(),
                             addColors.data(), addMaterialParams.data(), addFlags.data(), false);
                ++local;
            }
            nTraces += local;
        });
        selectClosest(hitPositions.data(), hitNormals.data(), hitColors.data(), hitMaterialParams.data(),
                      hitFlags.data(), nullptr);

        if (!lambertView) {
            initializePathState();
            const float* curPos = hitPositions.data();
            const float* curNrm = hitNormals.data();
            const float* curCol = hitColors.data();
            const float* curMat = hitMaterialParams.data();
            const int* curFlags = hitFlags.data();
            const float* curDist = hitDistances.data();
            const float* prevDirs = nullptr;
            for (int bounce = 1; bounce <= maxBounces; ++bounce) {
                sampleBounceDirections(curPos, curNrm, curCol, curMat, curFlags, curDist, bounceDistances.data(),
                                       prevDirs);
                forPixels(params, [&](int x, int y) {  // checkBounceEarlyTerminationKernel :503-553
                    int pixelIdx = y * params.width + x;
                    uint64_t local = 0;
                    for (int s = 0; s < params.samplesPerPixel; ++s) {
                        int sampleIdx = pixelIdx + s * params.pixelCount;
                        int base = sampleIdx * 3;
                        if (!pathActive[sampleIdx]) continue;
                        if (bouncePdfs[sampleIdx] <= 0.0f) continue;
                        if (bounceDistances[sampleIdx] <= 0.0f) continue;
                        Ray ray(get3(bounceOrigins.data(), base), get3(bounceDirections.data(), base));
                        HitInfo oh, ih;
                        bool hitOuter = traceMeshWithMode(ray, outerView, &oh, TraceMode::ANY, false, params.material);
                        bool hitInner = traceMeshWithMode(ray, innerView, &ih, TraceMode::ANY, false, params.material);
                        local += 2;
                        if (hitInner && hitOuter && (ih.t < oh.t)) pathActive[sampleIdx] = 0;
                    }
                    nTraces += local;
                });
                traceNeuralSegmentsForRays(false, bounceOrigins.data(), bounceDirections.data(), pathActive.data(),
                                           bouncePdfs.data(), elementCount, params, outerView, innerView, outerMin,
                                           outerInvExtent, bouncePositions.data(), bounceNormals.data(),
                                           bounceColors.data(), bounceMaterialParams.data(), bounceHitFlags.data(),
                                           bounceDistances.data());
                forPixels(params, [&](int x, int y) {  // traceAdditionalMeshRaysKernel :1235-1302
                    int pixelIdx = y * params.width + x;
                    uint64_t local = 0;
                    for (int s = 0; s < params.samplesPerPixel; ++s) {
                        int sampleIdx = pixelIdx + s * params.pixelCount;
                        int base = sampleIdx * 3;
                        if (additionalView.numTriangles == 0) { addFlags[sampleIdx] = 0; continue; }
                        if (bouncePdfs[sampleIdx] <= 0.0f) { addFlags[sampleIdx] = 0; continue; }
                        Ray ray(get3(bounceOrigins.data(), base), get3(bounceDirections.data(), base));
                        writeMeshHit(ray, additionalView, false, sampleIdx, addPositions.data(), addNormals.data(),
                                     addColors.data(), addMaterialParams.data(), addFlags.data(), false);
                        ++local;
                    }
                    nTraces += local;
                });
                selectClosest(bouncePositions.data(), bounceNormals.data(), bounceColors.data(),
                              bounceMaterialParams.data(), bounceHitFlags.data(), bounceOrigins.data());
                integrateBounce(bounce);
                prevBounceDirections = bounceDirections;  // D2D copy :2507
                prevDirs = prevBounceDirections.data();
                curPos = bouncePositions.data();
                curNrm = bounceNormals.data();
                curCol = bounceColors.data();
                curFlags = bounceHitFlags.data();
                curMat = bounceMaterialParams.data();
                curDist = bounceDistances.data();
            }
            finalize();
            if (denoise) denoisePass();
            accumSampleCount += static_cast<uint32_t>(spp);
        } else {
            lambert();
            accumSampleCount = 0;
        }
    } else {
        forPixels(params, [&](int x, int y) {  // intersectGroundTruthKernel :244-296
            int pixelIdx = y * params.width + x;
            for (int s = 0; s < params.samplesPerPixel; ++s) {
                int sampleIdx = pixelIdx + s * params.pixelCount;
                uint32_t rng = initRng(pixelIdx, params.sampleOffset, s);
                Ray ray = generatePrimaryRay(x, y, params, rng);
                writeMeshHit(ray, classicView, true, sampleIdx, hitPositions.data(), hitNormals.data(),
                             hitColors.data(), hitMaterialParams.data(), hitFlags.data(), true);
            }
            nTraces += static_cast<uint64_t>(params.samplesPerPixel);
        });
        if (lambertView) {
            lambert();
            accumSampleCount = 0;
        } else {
            initializePathState();
            const float* curPos = hitPositions.data();
            const float* curNrm = hitNormals.data();
            const float* curCol = hitColors.data();
            const float* curMat = hitMaterialParams.data();
            const int* curFlags = hitFlags.data();
            const float* prevDirs = nullptr;
            for (int bounce = 1; bounce <= maxBounces; ++bounce) {
                sampleBounceDirections(curPos, curNrm, curCol, curMat, curFlags, nullptr, nullptr, prevDirs);
                forPixels(params, [&](int x, int y) {  // traceGroundTruthBouncesKernel :556-618
                    int pixelIdx = y * params.width + x;
                    uint64_t local = 0;
                    for (int s = 0; s < params.samplesPerPixel; ++s) {
                        int sampleIdx = pixelIdx + s * params.pixelCount;
                        int base = sampleIdx * 3;
                        if (bouncePdfs[sampleIdx] <= 0.0f) {
                            bounceHitFlags[sampleIdx] = 0;
                            continue;
                        }
                        Ray ray(get3(bounceOrigins.data(), base), get3(bounceDirections.data(), base));
                        writeMeshHit(ray, classicView, false, sampleIdx, bouncePositions.data(),
                                     bounceNormals.data(), bounceColors.data(), bounceMaterialParams.data(),
                                     bounceHitFlags.data(), false);
                        ++local;
                    }
                    nTraces += local;
                });
                integrateBounce(bounce);
                prevBounceDirections = bounceDirections;
                prevDirs = prevBounceDirections.data();
                curPos = bouncePositions.data();
                curNrm = bounceNormals.data();
                curCol = bounceColors.data();
                curMat = bounceMaterialParams.data();
                curFlags = bounceHitFlags.data();
            }
            finalize();
            if (denoise) denoisePass();
            accumSampleCount += static_cast<uint32_t>(spp);
        }
    }
    counters[1] += nTraces.load();
    counters[3] += nVerts.load();
}

// ---------------------------------------------------------------------------------------------
// C entry points: same names and arguments as include/nrb200.h with the prefix orc_, so that the
// tests can drive product and oracle through one Python wrapper.
// ---------------------------------------------------------------------------------------------
static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}

extern "C" {

const char* orc_last_error(void) { return g_err.c_str(); }
const char* orc_backend(void) {
#ifdef NRB_ORACLE_REFERENCE
    return "reference";
#else
    return "port";
#endif
}

int orc_create(const nrb_nn_config* cfg, int /*device*/, orc_renderer** out) {
    if (!out) return fail(NRB_ERR_INVALID, "out is null");
    auto* r = new orc_renderer();
    r->net.configure(cfg);
    r->pool.threads = 1;
    {
        // the reference zero-fills its parameter block at construction and renders with it until a checkpoint is
        // loaded (cuda_renderer_neural.cu:1691-1697; neuralReady needs only the network and an outer shell, :2258)
        std::vector<uint16_t> zeros(r->net.fileBytes() / 2, 0);
        r->net.load(zeros.data(), true);
        r->weightsLoaded = true;
    }
    *out = r;
    return 0;
}
void orc_destroy(orc_renderer* r) { delete r; }
int orc_set_threads(orc_renderer* r, int n) {
    r->pool.threads = n <= 0 ? static_cast<int>(std::thread::hardware_concurrency()) : n;
    return r->pool.threads;
}
int orc_set_row_range(orc_renderer* r, int y0, int y1) {
    r->rowBegin = y0;
    r->rowEnd = y1;
    return 0;
}

// Visit bands of `band` rows, every `stride`-th band of the row range; returns the number of rows a frame of the
// given height then covers.
int orc_set_row_sample(orc_renderer* r, int band, int stride, int height) {
    r->rowBand = band < 1 ? 1 : band;
    r->rowStride = stride < 1 ? 1 : stride;
    int y0 = std::max(0, r->rowBegin), y1 = std::min(height, r->rowEnd), n = 0;
    for (int y = y0; y < y1; ++y)
        if (r->rowStride <= 1 || ((y - y0) / r->rowBand) % r->rowStride == 0) ++n;
    return n;
}

int orc_update_mesh(orc_renderer* r, int slot, const nrb_mesh_desc* d) {
    if (slot < 0 || slot >= 4) return fail(NRB_ERR_INVALID, "bad slot");
    OrcMesh& m = r->mesh[slot];
    m = OrcMesh();
    if (!d || d->num_triangles <= 0) return 0;
    const Vec3* v = reinterpret_cast<const Vec3*>(d->vertices);
    m.vertices.assign(v, v + d->num_vertices);
    if (d->normals) {
        const Vec3* n = reinterpret_cast<const Vec3*>(d->normals);
        m.normals.assign(n, n + d->num_vertices);
    }
    if (d->texcoords) {
        const Vec2* t = reinterpret_cast<const Vec2*>(d->texcoords);
        m.texcoords.assign(t, t + d->num_vertices);
    }
    m.indices.resize(d->num_triangles);
    std::memcpy(m.indices.data(), d->indices, sizeof(uint32_t) * 3 * d->num_triangles);
    m.nodes.resize(d->num_nodes);
    std::memcpy(m.nodes.data(), d->nodes, sizeof(BvhNode) * d->num_nodes);
    if (d->materials && d->num_materials > 0) {
        m.materials.resize(d->num_materials);
        std::memcpy(static_cast<void*>(m.materials.data()), d->materials, sizeof(Material) * d->num_materials);
    }
    if (d->material_map && d->material_ids && d->num_primitives > 0) {
        m.materialMap.assign(d->material_map, d->material_map + d->num_primitives);
        m.materialIds.assign(d->material_ids, d->material_ids + d->num_primitives);
    }
    for (int i = 0; i < d->num_textures; ++i) {
        const nrb_texture& t = d->textures[i];
        size_t bytes = static_cast<size_t>(t.width) * t.height * t.channels;
        m.texPixels.emplace_back(t.pixels, t.pixels + bytes);
    }
    for (int i = 0; i < d->num_textures; ++i) {
        TextureDeviceView tv{};
        tv.pixels = m.texPixels[i].data();
        tv.width = d->textures[i].width;
        tv.height = d->textures[i].height;
        tv.channels = d->textures[i].channels;
        tv.srgb = d->textures[i].srgb;
        m.textures.push_back(tv);
    }
    m.textureNearest = d->texture_nearest;
    m.hasMeshMaterials = d->has_mesh_materials;
    m.bmin = Vec3(d->bounds_min[0], d->bounds_min[1], d->bounds_min[2]);
    m.bmax = Vec3(d->bounds_max[0], d->bounds_max[1], d->bounds_max[2]);
    return 0;
}

int orc_update_env(orc_renderer* r, const float* rgb, int w, int h, float strength) {
    r->envPixels.clear();
    r->envW = r->envH = 0;
    r->envStrength = strength;
    if (rgb && w > 0 && h > 0) {
        const Vec3* p = reinterpret_cast<const Vec3*>(rgb);
        r->envPixels.assign(p, p + static_cast<size_t>(w) * h);
        r->envW = w;
        r->envH = h;
    }
    return 0;
}
int orc_set_global_material(orc_renderer* r, const nrb_material* m) {
    std::memcpy(static_cast<void*>(&r->globalMaterial), m, sizeof(Material));
    return 0;
}
int orc_load_weights(orc_renderer* r, const void* fp16, size_t bytes, int swap) {
    if (bytes != r->net.fileBytes()) return fail(NRB_ERR_SIZE, "weights size mismatch");
    r->net.load(static_cast<const uint16_t*>(fp16), swap != 0);
    r->weightsLoaded = true;
    return 0;
}
size_t orc_weights_file_bytes(const orc_renderer* r) { return r->net.fileBytes(); }
size_t orc_params_bytes(const orc_renderer* r) { return r->net.paramsBytes(); }
int orc_resize(orc_renderer* r, int w, int h) {
    if (w == r->width && h == r->height) return 0;
    r->width = w;
    r->height = h;
    r->bufferElements = 0;
    r->accum.clear();
    r->pixels.clear();
    return 0;
}
int orc_set_camera_basis(orc_renderer* r, const float f[3], const float rt[3], const float up[3], float fovY) {
    r->basis.forward = Vec3(f[0], f[1], f[2]);
    r->basis.right = Vec3(rt[0], rt[1], rt[2]);
    r->basis.up = Vec3(up[0], up[1], up[2]);
    r->basis.fovY = fovY;
    return 0;
}
int orc_set_option(orc_renderer* r, int opt, int v) {
    switch (opt) {
        case NRB_OPT_USE_NEURAL_QUERY: r->useNeuralQuery = v != 0; break;
        case NRB_OPT_SAMPLES_PER_PIXEL: r->samplesPerPixel = v; break;
        case NRB_OPT_BOUNCE_COUNT: r->bounceCount = v; break;
        case NRB_OPT_LAMBERT_VIEW: r->lambertView = v != 0; break;
        case NRB_OPT_CLASSIC_MESH_INDEX: r->classicMeshIndex = v; break;
        case NRB_OPT_USE_ADDITIONAL_MESH: r->useAdditionalMesh = v != 0; break;
        case NRB_OPT_USE_CONSTANT_NEURAL_COLOR: r->useConstantNeuralColor = v != 0; break;
        case NRB_OPT_USE_DIRECT_ENV_COLOR: r->useDirectEnvColor = v != 0; break;
        case NRB_OPT_DENOISE: r->denoise = v != 0; break;
        case NRB_OPT_CAPTURE: case NRB_OPT_USE_HARDWARE_RT: case NRB_OPT_LITERAL_BOUNCE_TRACES: break;
        default: return fail(NRB_ERR_INVALID, "unknown option");
    }
    return 0;
}
int orc_set_envmap_rotation(orc_renderer* r, float deg) { r->envmapRotation = deg; return 0; }
int orc_set_constant_neural_color(orc_renderer* r, const float c[3]) {
    r->constantNeuralColor = Vec3(c[0], c[1], c[2]);
    return 0;
}
int orc_set_direct_env_color(orc_renderer* r, const float c[3]) {
    r->directEnvColor = Vec3(c[0], c[1], c[2]);
    return 0;
}
int orc_reset_samples(orc_renderer* r) { r->resetAccum(); return 0; }
int orc_render(orc_renderer* r, const float cam[3]) {
    r->render(Vec3(cam[0], cam[1], cam[2]));
    return 0;
}
int orc_read_pixels(orc_renderer* r, void* rgba, size_t bytes) {
    if (bytes != r->pixels.size() * 4) return fail(NRB_ERR_SIZE, "pixel buffer size");
    std::memcpy(rgba, r->pixels.data(), bytes);
    return 0;
}
int orc_width(const orc_renderer* r) { return r->width; }
int orc_height(const orc_renderer* r) { return r->height; }
uint32_t orc_accum_sample_count(const orc_renderer* r) { return r->accumSampleCount; }
int orc_get_counters(orc_renderer* r, uint64_t out[4]) {
    for (int i = 0; i < 4; ++i) out[i] = r->counters[i];
    return 0;
}

int orc_read_buffer(orc_renderer* r, const char* name, void* dst, size_t bytes) {
    std::string n(name);
    size_t elems = static_cast<size_t>(r->width) * r->height * std::max(1, r->samplesPerPixel);
    const void* src = nullptr;
    size_t have = 0;
    auto f = [&](const std::vector<float>& v, size_t per) { src = v.data(); have = elems * per * 4; };
    if (n == "hit_positions") f(r->hitPositions, 3);
    else if (n == "hit_normals") f(r->hitNormals, 3);
    else if (n == "hit_colors") f(r->hitColors, 3);
    else if (n == "hit_material_params") f(r->hitMaterialParams, 3);
    else if (n == "hit_distances") f(r->hitDistances, 1);
    else if (n == "hit_flags") { src = r->hitFlags.data(); have = elems * 4; }
    else if (n == "path_radiance") { src = r->pathRadiance.data(); have = elems * 12; }
    else if (n == "accum") { src = r->accum.data(); have = r->accum.size() * 12; }
    else return fail(NRB_ERR_INVALID, "unknown buffer " + n);
    if (bytes != have) return fail(NRB_ERR_SIZE, "buffer size mismatch for " + n);
    std::memcpy(dst, src, bytes);
    return 0;
}

// traceMeshWithMode on caller rays; mirrors nrb_trace_rays_host.  PreliminaryHitInfo is
// recovered from HitInfo.t plus a re-run of the triangle test for (u, v): the reference
// function only returns HitInfo, so (t,u,v,prim) are taken from an identical traversal
// restated in trace_prelim() below.
int orc_trace_rays_host(orc_renderer* r, int slot, int mode, const float* origins, const float* dirs, size_t n,
                        nrb_hit* out);
int orc_neural_query_host(orc_renderer* r, const float* inputs, size_t n, void* out_fp16) {
    if (!r->weightsLoaded) return fail(NRB_ERR_STATE, "no weights");
    const int D = r->net.inputDims();
    r->pool.run(static_cast<int>(n), [&](int a, int b) {
        r->net.forward(inputs + static_cast<size_t>(a) * D, b - a, static_cast<uint16_t*>(out_fp16) + static_cast<size_t>(a) * 16);
    });
    return 0;
}
int orc_input_dims(const orc_renderer* r) { return r->net.inputDims(); }
int orc_output_dims(const orc_renderer*) { return 16; }

}  // extern "C"

// traceMeshWithMode only returns HitInfo.  (u, v, prim) are read back through it by tracing a
// view without texcoords (then HitInfo.uv = (u, v), mesh_intersection.cpp:204-206) and with an
// identity material map (then HitInfo.materialId = triangle index, :53-85) and no materials
// (so the normal-map branch :212-235 is skipped).  No traversal code is restated for this.
extern "C" int orc_trace_rays_host(orc_renderer* r, int slot, int mode, const float* origins, const float* dirs,
                                   size_t n, nrb_hit* out) {
    if (slot < 0 || slot >= 4) return fail(NRB_ERR_INVALID, "bad slot");
    const OrcMesh& m = r->mesh[slot];
    MeshDeviceView v = m.view();
    std::vector<uint32_t> ident(m.indices.size());
    std::vector<int> identI(m.indices.size());
    for (size_t i = 0; i < ident.size(); ++i) {
        ident[i] = static_cast<uint32_t>(i);
        identI[i] = static_cast<int>(i);
    }
    v.texcoords = nullptr;
    v.hasTexcoords = 0;
    v.materials = nullptr;
    v.numMaterials = 0;
    v.materialMap = ident.data();
    v.materialIds = identI.data();
    v.numPrimitives = static_cast<int>(ident.size());
    v.hasMeshMaterials = 1;
    TraceMode tm = mode == NRB_TRACE_FORWARD_ONLY ? TraceMode::FORWARD_ONLY
                   : mode == NRB_TRACE_BACKWARD_ONLY ? TraceMode::BACKWARD_ONLY
                                                     : TraceMode::ANY;
    Material g = r->globalMaterial;
    r->pool.run(static_cast<int>(n), [&](int a, int b) {
        for (int i = a; i < b; ++i) {
            Ray ray(Vec3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
                    Vec3(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
            HitInfo h;
            bool hit = traceMeshWithMode(ray, v, &h, tm, false, g);
            if (hit) {
                out[i].t = h.t;
                out[i].u = h.uv.x;
                out[i].v = h.uv.y;
                out[i].prim = static_cast<uint32_t>(h.materialId);
            } else {
                out[i].t = 1e30f;
                out[i].u = 0.0f;
                out[i].v = 0.0f;
                out[i].prim = 0;
            }
        }
    });
    return 0;
}
