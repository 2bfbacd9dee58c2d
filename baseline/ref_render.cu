// baseline/ref_render.cu - headless driver of the REFERENCE's own renderer (RendererNeural, src/cuda_renderer_neural.cu)
// + its own Mesh::buildBvh (src/mesh_bvh.cpp) + tiny-cuda-nn, compiled for sm_100 from /root/reference by
// baseline/Makefile.  It is the GPU arbiter and the timed GPU baseline of SURVEY.md 8c row 2 / 8d:
//   * renders a scene pack (baseline/refpack.py: the same seeded synthetic scenes, camera, weights the product gets)
//     through the reference's public API in evaluate's call order (src/evaluate.cu:211-296),
//   * dumps the BVH the reference's builder produced (nodes + leaf-ordered indices = the primitive ids), the primary
//     hit buffers, path radiance, accumulation and the 8-bit frame,
//   * times `timed` render() calls after `warmup` with cudaEvents on the legacy default stream (which is where the
//     reference enqueues everything), and prints one JSON line.
// Private members are read through `#define private public` in THIS translation unit only (layout is unaffected);
// the reference's own TUs are compiled untouched.  Test/bench infrastructure: never linked by libnrb200.
//
//   ref_render <scene.pack> <out_dir> [warmup timed]
#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#define private public
#include "scene.h"
#include "cuda_renderer_neural.h"
#undef private
#include "config_loader.h"

// scene.cu references the EXR loader; ext/tinyexr is an empty (un-vendored) submodule here and the harness never
// loads EXR files, so the two symbols are satisfied by failing stubs.
extern "C" {
int LoadEXR(float**, int*, int*, const char*, const char** err) {
    if (err) *err = nullptr;
    return -1;
}
void FreeEXRErrorMessage(const char*) {}
}

namespace {

struct Reader {
    std::vector<char> buf;
    size_t pos = 0;
    explicit Reader(const char* path) {
        std::ifstream f(path, std::ios::binary | std::ios::ate);
        if (!f) { std::fprintf(stderr, "cannot open %s\n", path); std::exit(2); }
        buf.resize(static_cast<size_t>(f.tellg()));
        f.seekg(0);
        f.read(buf.data(), static_cast<std::streamsize>(buf.size()));
    }
    void raw(void* dst, size_t n) {
        if (pos + n > buf.size()) { std::fprintf(stderr, "scene pack truncated\n"); std::exit(2); }
        std::memcpy(dst, buf.data() + pos, n);
        pos += n;
    }
    template <class T> T get() { T v; raw(&v, sizeof(T)); return v; }
    template <class T> void vec(std::vector<T>& v, size_t n) { v.resize(n); if (n) raw(v.data(), n * sizeof(T)); }
};

struct Params {
    int32_t width, height, spp, bounces, neural, lambert, classicMeshIndex, useAdditional, useConstColor, useDirectEnv,
        log2T, baseRes, accumFrames, swapOrder, reserved[2];
    float camPos[3], fwd[3], right[3], up[3], fovY, envRotation, envStrength, constColor[3], directEnv[3], pad[3];
};
static_assert(sizeof(Params) == 16 * 4 + 24 * 4, "pack params layout");

void dump(const std::string& path, const void* p, size_t bytes) {
    std::ofstream f(path, std::ios::binary);
    f.write(static_cast<const char*>(p), static_cast<std::streamsize>(bytes));
}
void dumpDev(const std::string& path, const void* dev, size_t bytes) {
    if (!dev) return;
    std::vector<char> h(bytes);
    cudaMemcpy(h.data(), dev, bytes, cudaMemcpyDeviceToHost);
    dump(path, h.data(), bytes);
}

void loadMesh(Reader& r, Mesh& m) {
    const int32_t nv = r.get<int32_t>(), nt = r.get<int32_t>(), hasN = r.get<int32_t>(), hasUv = r.get<int32_t>(),
                  nNodes = r.get<int32_t>(), nMat = r.get<int32_t>(), nPrim = r.get<int32_t>(), nTex = r.get<int32_t>(),
                  texNearest = r.get<int32_t>(), hasMeshMat = r.get<int32_t>();
    r.vec(m.vertices_, nv);
    if (hasN) r.vec(m.normals_, nv);
    if (hasUv) r.vec(m.texcoords_, nv);
    r.vec(m.indices_, nt);
    std::vector<BvhNode> nodes;
    r.vec(nodes, nNodes);
    static_assert(sizeof(Material) == 180 && sizeof(BvhNode) == 44, "reference struct sizes");
    r.vec(m.materials_, nMat);
    r.vec(m.materialMap_, nPrim);
    r.vec(m.materialIds_, nPrim);
    m.textures_.resize(nTex);
    for (int i = 0; i < nTex; ++i) {
        MeshTexture& t = m.textures_[i];
        t.width = r.get<int32_t>(); t.height = r.get<int32_t>(); t.channels = r.get<int32_t>();
        t.colorSpace = r.get<int32_t>() ? ColorSpace::SRGB : ColorSpace::LINEAR;
        r.vec(t.pixels, static_cast<size_t>(t.width) * t.height * t.channels);
    }
    m.textureNearest_ = texNearest != 0;
    m.hasMeshMaterials_ = hasMeshMat != 0;
    if (nt == 0) return;
    if (nNodes > 0) {
        // a BVH supplied by the pack (indices already in its leaf order): install it as buildBvh would have left it
        m.bvhNodes_ = nodes;
        Vec3 mn(FLT_MAX, FLT_MAX, FLT_MAX), mx(-FLT_MAX, -FLT_MAX, -FLT_MAX);
        for (const Vec3& v : m.vertices_) {
            mn.x = fminf(mn.x, v.x); mn.y = fminf(mn.y, v.y); mn.z = fminf(mn.z, v.z);
            mx.x = fmaxf(mx.x, v.x); mx.y = fmaxf(mx.y, v.y); mx.z = fmaxf(mx.z, v.z);
        }
        m.boundsMin_ = mn; m.boundsMax_ = mx;
        m.bvhDirty_ = false; m.boundsDirty_ = false;
    } else {
        m.buildBvh();  // the reference's own builder: defines leaf order = primitive ids
    }
}

void dumpBvh(const std::string& path, const Mesh& m) {
    std::ofstream f(path, std::ios::binary);
    const int32_t hdr[4] = {static_cast<int32_t>(m.bvhNodes_.size()), m.numTriangles(), static_cast<int32_t>(m.materialMap_.size()), 0};
    f.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
    f.write(reinterpret_cast<const char*>(m.bvhNodes_.data()), static_cast<std::streamsize>(m.bvhNodes_.size() * sizeof(BvhNode)));
    f.write(reinterpret_cast<const char*>(m.indices_.data()), static_cast<std::streamsize>(m.indices_.size() * sizeof(uint3)));
    f.write(reinterpret_cast<const char*>(m.materialMap_.data()), static_cast<std::streamsize>(m.materialMap_.size() * 4));
    f.write(reinterpret_cast<const char*>(m.materialIds_.data()), static_cast<std::streamsize>(m.materialIds_.size() * 4));
}

}  // namespace

int main(int argc, char** argv) {
    if (argc < 3) {
        std::fprintf(stderr, "usage: ref_render <scene.pack> <out_dir> [warmup timed]\n");
        return 2;
    }
    const std::string outDir = argv[2];
    const int warmup = argc > 3 ? std::atoi(argv[3]) : 0, timed = argc > 4 ? std::atoi(argv[4]) : 0;
    Reader r(argv[1]);
    char magic[8];
    r.raw(magic, 8);
    if (std::memcmp(magic, "NRBPACK1", 8) != 0) { std::fprintf(stderr, "bad scene pack magic\n"); return 2; }
    const Params p = r.get<Params>();

    Scene scene;
    Mesh* meshes[4] = {&scene.originalMesh(), &scene.innerShell(), &scene.outerShell(), &scene.additionalMesh()};
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < 4; ++i) loadMesh(r, *meshes[i]);
    const double buildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    {
        const int32_t w = r.get<int32_t>(), h = r.get<int32_t>();
        EnvironmentMap& env = scene.environment();
        if (w > 0 && h > 0) {
            r.vec(env.pixels_, static_cast<size_t>(w) * h);
            env.width_ = w; env.height_ = h; env.deviceDirty_ = true;
        }
        env.setRotation(p.envRotation);
        env.setStrength(p.envStrength);
    }
    r.raw(&scene.globalMaterial(), sizeof(Material));
    const int64_t wbytes = r.get<int64_t>();
    std::vector<char> weights;
    r.vec(weights, static_cast<size_t>(wbytes));
    for (int i = 0; i < 4; ++i)
        if (meshes[i]->numTriangles() > 0) dumpBvh(outDir + "/bvh_" + std::to_string(i) + ".bin", *meshes[i]);

    // evaluate.cu:211-232
    NeuralNetworkConfig nn;
    nn.log2_hashmap_size = p.log2T;
    nn.base_resolution = p.baseRes;
    nn.use_neural_query = p.neural != 0;
    RendererNeural renderer(scene, &nn);
    renderer.setConstantNeuralColor(p.useConstColor != 0, Vec3(p.constColor[0], p.constColor[1], p.constColor[2]));
    renderer.setDirectEnvColor(p.useDirectEnv != 0, Vec3(p.directEnv[0], p.directEnv[1], p.directEnv[2]));
    renderer.resize(p.width, p.height);
    renderer.setBounceCount(p.bounces);
    renderer.setLambertView(p.lambert != 0);
    renderer.setEnvmapRotation(p.envRotation);
    renderer.setUseAdditionalMesh(p.useAdditional != 0);
    renderer.setClassicMeshIndex(p.classicMeshIndex);
    renderer.setSwapParamOrder(p.swapOrder != 0);
    if (wbytes > 0) {
        const std::string wpath = outDir + "/_weights.bin";
        dump(wpath, weights.data(), weights.size());
        if (!renderer.loadWeightsFromFile(wpath)) return 3;
    }
    renderer.setUseNeuralQuery(p.neural != 0);
    RenderBasis basis;
    basis.forward = Vec3(p.fwd[0], p.fwd[1], p.fwd[2]);
    basis.right = Vec3(p.right[0], p.right[1], p.right[2]);
    basis.up = Vec3(p.up[0], p.up[1], p.up[2]);
    basis.fovY = p.fovY;
    renderer.setCameraBasis(basis);
    renderer.setSamplesPerPixel(p.spp);
    renderer.resetSamples();
    const Vec3 camPos(p.camPos[0], p.camPos[1], p.camPos[2]);

    // the frames to dump: accumFrames render() calls (progressive accumulation), buffers of the last one
    for (int i = 0; i < (p.accumFrames > 0 ? p.accumFrames : 1); ++i) renderer.render(camPos);
    cudaDeviceSynchronize();
    const size_t px = static_cast<size_t>(p.width) * p.height, n = px * static_cast<size_t>(p.spp > 0 ? p.spp : 1);
    dumpDev(outDir + "/pixels.rgba", renderer.devicePixels(), px * 4);
    dumpDev(outDir + "/hit_flags.i32", renderer.hitFlags_, n * 4);
    dumpDev(outDir + "/hit_positions.f32", renderer.hitPositions_, n * 12);
    dumpDev(outDir + "/hit_normals.f32", renderer.hitNormals_, n * 12);
    dumpDev(outDir + "/hit_colors.f32", renderer.hitColors_, n * 12);
    dumpDev(outDir + "/hit_material_params.f32", renderer.hitMaterialParams_, n * 12);
    dumpDev(outDir + "/hit_distances.f32", renderer.hitDistances_, n * 4);
    dumpDev(outDir + "/path_radiance.f32", renderer.pathRadiance_, n * 12);
    dumpDev(outDir + "/accum.f32", renderer.accum_, px * 12);

    double msPerFrame = 0.0;
    std::string buckets;
    if (timed > 0) {
        cudaEvent_t a, b;
        cudaEventCreate(&a);
        cudaEventCreate(&b);
        for (int i = 0; i < warmup; ++i) renderer.render(camPos);
        cudaDeviceSynchronize();
#ifdef PROFILE_KERNELS
        double sum[KID_COUNT] = {};
        uint64_t calls = 0;
#endif
        const auto h0 = std::chrono::steady_clock::now();
        cudaEventRecord(a, 0);
        for (int i = 0; i < timed; ++i) {
            renderer.render(camPos);
#ifdef PROFILE_KERNELS
            const KernelTimings& t = renderer.lastFrameTimings();
            for (int k = 0; k < KID_COUNT; ++k) sum[k] += t.ms[k];
            calls += t.neuralRayCalls;
#endif
        }
        cudaEventRecord(b, 0);
        cudaEventSynchronize(b);
        const double hostMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - h0).count();
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, a, b);
        msPerFrame = ms / timed;
        char tmp[256];
        std::snprintf(tmp, sizeof(tmp), ", \"host_ms_per_frame\": %.4f", hostMs / timed);
        buckets += tmp;
#ifdef PROFILE_KERNELS
        buckets += ", \"buckets_ms\": {";
        for (int k = 0; k < KID_COUNT; ++k) {
            std::snprintf(tmp, sizeof(tmp), "%s\"%s\": %.4f", k ? ", " : "", KernelTimings::name(k), sum[k] / timed);
            buckets += tmp;
        }
        std::snprintf(tmp, sizeof(tmp), "}, \"neural_ray_calls\": %.1f", static_cast<double>(calls) / timed);
        buckets += tmp;
#endif
    }
    const double rays = static_cast<double>(n);
    char line[4096];
    std::snprintf(line, sizeof(line),
                  "{\"impl\": \"reference-cuda\", \"mode\": \"%s\", \"width\": %d, \"height\": %d, \"spp\": %d, \"bounces\": %d, "
                  "\"warmup\": %d, \"timed\": %d, \"ms_per_frame\": %.4f, \"mrays_per_s\": %.2f, \"bvh_build_ms\": %.2f, "
                  "\"tris\": [%d, %d, %d, %d], \"profile_kernels\": %s%s}",
                  p.neural ? "neural" : "gt", p.width, p.height, p.spp, p.bounces, warmup, timed, msPerFrame,
                  msPerFrame > 0 ? rays / msPerFrame * 1e-3 : 0.0, buildMs, meshes[0]->numTriangles(), meshes[1]->numTriangles(),
                  meshes[2]->numTriangles(), meshes[3]->numTriangles(),
#ifdef PROFILE_KERNELS
                  "true",
#else
                  "false",
#endif
                  buckets.c_str());
    std::printf("%s\n", line);
    dump(outDir + "/timing.json", line, std::strlen(line));
    return 0;
}
