// baseline/pack_reader.h - scene-pack reader shared by the two harnesses (ref_render.cu: the reference's own renderer;
// shim_driver.cu: include/nrb200_renderer.hpp over libnrb200.so).  Fills the reference's OWN Scene / Mesh /
// EnvironmentMap objects from the flat file baseline/refpack.py writes.  Private members are reached through
// `#define private public` in the including translation unit only (layout is unaffected); test infrastructure.
#pragma once
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
#undef private
#include "config_loader.h"

// scene.cu references the EXR loader; ext/tinyexr is an empty (un-vendored) submodule here and the harness never
// loads EXR files, so the two symbols are satisfied by failing stubs.
#ifdef PACK_DEFINE_EXR_STUBS
extern "C" {
int LoadEXR(float**, int*, int*, const char*, const char** err) {
    if (err) *err = nullptr;
    return -1;
}
void FreeEXRErrorMessage(const char*) {}
}
#endif

namespace pack {

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

inline void dump(const std::string& path, const void* p, size_t bytes) {
    std::ofstream f(path, std::ios::binary);
    f.write(static_cast<const char*>(p), static_cast<std::streamsize>(bytes));
}
inline void dumpDev(const std::string& path, const void* dev, size_t bytes) {
    if (!dev) return;
    std::vector<char> h(bytes);
    cudaMemcpy(h.data(), dev, bytes, cudaMemcpyDeviceToHost);
    dump(path, h.data(), bytes);
}

inline void loadMesh(Reader& r, Mesh& m) {
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

inline void dumpBvh(const std::string& path, const Mesh& m) {
    std::ofstream f(path, std::ios::binary);
    const int32_t hdr[4] = {static_cast<int32_t>(m.bvhNodes_.size()), m.numTriangles(), static_cast<int32_t>(m.materialMap_.size()), 0};
    f.write(reinterpret_cast<const char*>(hdr), sizeof(hdr));
    f.write(reinterpret_cast<const char*>(m.bvhNodes_.data()), static_cast<std::streamsize>(m.bvhNodes_.size() * sizeof(BvhNode)));
    f.write(reinterpret_cast<const char*>(m.indices_.data()), static_cast<std::streamsize>(m.indices_.size() * sizeof(uint3)));
    f.write(reinterpret_cast<const char*>(m.materialMap_.data()), static_cast<std::streamsize>(m.materialMap_.size() * 4));
    f.write(reinterpret_cast<const char*>(m.materialIds_.data()), static_cast<std::streamsize>(m.materialIds_.size() * 4));
}


// Read the four meshes (running the reference's Mesh::buildBvh unless the pack carries nodes), the environment, the
// global material and the weight blob; returns the milliseconds spent in the mesh part.
inline double loadScene(Reader& r, const Params& p, Scene& scene, std::vector<char>& weights) {
    Mesh* meshes[4] = {&scene.originalMesh(), &scene.innerShell(), &scene.outerShell(), &scene.additionalMesh()};
    const auto t0 = std::chrono::steady_clock::now();
    for (int i = 0; i < 4; ++i) loadMesh(r, *meshes[i]);
    const double buildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    const int32_t w = r.get<int32_t>(), h = r.get<int32_t>();
    EnvironmentMap& env = scene.environment();
    if (w > 0 && h > 0) {
        r.vec(env.pixels_, static_cast<size_t>(w) * h);
        env.width_ = w; env.height_ = h; env.deviceDirty_ = true;
    }
    env.setRotation(p.envRotation);
    env.setStrength(p.envStrength);
    r.raw(&scene.globalMaterial(), sizeof(Material));
    const int64_t wbytes = r.get<int64_t>();
    r.vec(weights, static_cast<size_t>(wbytes));
    return buildMs;
}

inline Params openPack(Reader& r) {
    char magic[8];
    r.raw(magic, 8);
    if (std::memcmp(magic, "NRBPACK1", 8) != 0) { std::fprintf(stderr, "bad scene pack magic\n"); std::exit(2); }
    return r.get<Params>();
}

}  // namespace pack
