// oracle/restated/geometry.h — TEST INFRASTRUCTURE (CPU oracle "port" build), never part of the product.
//
// Plain C++ restatement of the reference's per-ray geometry, used when the reference's own headers
// are not available (the GPU box).  Types keep the reference's names and member layouts because
// frame_oracle.cpp is compiled against either this file or the real headers (backend.h).  Each
// block cites the reference lines it follows.  Pinned against the reference headers themselves by
// tests/test_oracle_pinning.py (both builds run on the same inputs wherever /root/reference exists).
#pragma once

#include <cmath>
#include <cstdint>

// ---- src/vec3.h ---------------------------------------------------------------------------------
struct Vec3 {
    float x = 0.0f, y = 0.0f, z = 0.0f;
    Vec3() = default;
    Vec3(float a, float b, float c) : x(a), y(b), z(c) {}
};
struct Vec2 {
    float x = 0.0f, y = 0.0f;
    Vec2() = default;
    Vec2(float a, float b) : x(a), y(b) {}
};
inline Vec2 operator+(Vec2 a, Vec2 b) { return {a.x + b.x, a.y + b.y}; }
inline Vec2 operator*(Vec2 a, float s) { return {a.x * s, a.y * s}; }
inline Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator-(Vec3 a) { return {-a.x, -a.y, -a.z}; }
inline Vec3 operator*(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline Vec3 operator*(float s, Vec3 a) { return a * s; }
inline Vec3 operator/(Vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline Vec3& operator+=(Vec3& a, Vec3 b) { return a = a + b; }
inline float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float length(Vec3 v) { return sqrtf(dot(v, v)); }
inline Vec3 normalize(Vec3 v) {
    const float l = length(v);
    return l > 0.0f ? v / l : Vec3();
}
inline Vec3 lerp(Vec3 a, Vec3 b, float t) { return a * (1.0f - t) + b * t; }

struct uint3 {
    unsigned int x, y, z;
};

// ---- src/ray.h, src/hit_info.h, src/bvh_data.h ----------------------------------------------------
struct Ray {
    Vec3 origin, direction;
    Ray() = default;
    Ray(Vec3 o, Vec3 d) : origin(o), direction(d) {}
    Vec3 at(float t) const { return origin + direction * t; }
};
struct PreliminaryHitInfo {
    float t = 1e30f, u = 0.0f, v = 0.0f;
    uint32_t primIdx = 0;
    bool valid() const { return t < 1e30f; }
};
struct HitInfo {
    Vec3 position;
    float t = 1e30f;
    Vec3 shadingNormal, geometricNormal;
    Vec2 uv;
    int materialId = -1;
    bool valid() const { return t < 1e30f; }
};
struct BvhNode {
    Vec3 boundsMin, boundsMax;
    int left, right, first, count, isLeaf;
};

// ---- src/material.h:9-117 ---------------------------------------------------------------------------
struct MaterialParam {
    float value = 0.0f;
    uint32_t textureId = 0;
    uint8_t channel = 0;
    bool textured = false;
    static MaterialParam constant(float v) {
        MaterialParam p;
        p.value = v;
        return p;
    }
};
struct MaterialParamVec3 {
    Vec3 value;
    uint32_t textureId = 0;
    bool textured = false;
    static MaterialParamVec3 constant(Vec3 v) {
        MaterialParamVec3 p;
        p.value = v;
        return p;
    }
};
struct Material {
    MaterialParamVec3 base_color;
    MaterialParam metallic, roughness, specular, specular_tint, anisotropy, sheen, sheen_tint, clearcoat, clearcoat_gloss;
    MaterialParamVec3 normal, base_emission;
    float emission_scale = 1.0f, ior = 1.5f, specular_transmission = 0.0f;
    static Material defaultMaterial() {  // :86-106
        Material m;
        m.base_color = MaterialParamVec3::constant(Vec3(0.8f, 0.8f, 0.8f));
        m.clearcoat_gloss = MaterialParam::constant(1.0f);
        return m;
    }
};
static_assert(sizeof(MaterialParam) == 12 && sizeof(MaterialParamVec3) == 20 && sizeof(Material) == 180, "material layout");

// ---- src/mesh.h:31-78, src/scene.h:10-16, src/renderer.h:5-10, src/render_params.h -----------------
struct TextureDeviceView {
    const unsigned char* pixels = nullptr;
    int width = 0, height = 0, channels = 0, srgb = 0;
};
struct MeshDeviceView {
    const Vec3* vertices = nullptr;
    const Vec3* normals = nullptr;
    const Vec2* texcoords = nullptr;
    const uint3* indices = nullptr;
    int numTriangles = 0, numVertices = 0;
    const BvhNode* bvhNodes = nullptr;
    int numBvhNodes = 0;
    const Material* materials = nullptr;
    const uint32_t* materialMap = nullptr;
    const int* materialIds = nullptr;
    int numPrimitives = 0, numMaterials = 0;
    const TextureDeviceView* textures = nullptr;
    int numTextures = 0, textureNearest = 0;
    int hasMeshMaterials = 0, hasNormals = 0, hasTexcoords = 0;
};
struct EnvironmentDeviceView {
    const Vec3* pixels = nullptr;
    int width = 0, height = 0;
    float rotation = 0.0f, strength = 1.0f;
};
struct RenderBasis {
    Vec3 forward, right, up;
    float fovY = 0.0f;
};
struct RenderParams {
    Vec3 camPos, camForward, camRight, camUp, lightDir, outerShellMin, outerShellInvExtent;
    Material material;
    bool useConstantNeuralColor = false;
    Vec3 constantNeuralColor;
    bool useDirectEnvColor = false;
    Vec3 directEnvColor;
    float fovY = 0.0f, maxRadiance = 0.0f, sceneScale = 1.0f;
    int maxBounces = 0, width = 0, height = 0, pixelCount = 0, samplesPerPixel = 1;
    uint32_t sampleOffset = 0;
};

// ---- RNG and primary rays: src/mesh_traversal.cuh:17-55 --------------------------------------------
inline uint32_t wangHash(uint32_t v) {
    v = (v ^ 61u) ^ (v >> 16);
    v *= 9u;
    v ^= v >> 4;
    v *= 0x27d4eb2du;
    v ^= v >> 15;
    return v;
}
inline uint32_t initRng(int pixelIdx, uint32_t sampleOffset, int sampleIdx) {
    return wangHash(static_cast<uint32_t>(pixelIdx) * 9781u + (sampleOffset + sampleIdx + 1u) * 6271u);
}
inline float rand01(uint32_t& state) {
    state = wangHash(state);
    return (state & 0x00FFFFFFu) * (1.0f / 16777216.0f);
}
inline Ray generatePrimaryRay(int x, int y, const RenderParams& p, uint32_t& rng) {
    const float jx = rand01(rng), jy = rand01(rng);
    const float aspect = static_cast<float>(p.width) / static_cast<float>(p.height);
    float u = (static_cast<float>(x) + jx) / static_cast<float>(p.width);
    float v = 1.0f - (static_cast<float>(y) + jy) / static_cast<float>(p.height);
    u = u * 2.0f - 1.0f;
    v = v * 2.0f - 1.0f;
    const float th = tanf(p.fovY * 0.5f);
    return Ray(p.camPos, normalize(p.camForward + p.camRight * (u * aspect * th) + p.camUp * (v * th)));
}

// ---- slab test: src/mesh_traversal.cuh:61-93 --------------------------------------------------------
inline bool intersectAabb(const Ray& ray, const Vec3& inv, const Vec3& bmin, const Vec3& bmax, float tMax, float* tNear) {
    const float pad = 1e-10f;
    const float lo[3] = {bmin.x - pad, bmin.y - pad, bmin.z - pad}, hi[3] = {bmax.x + pad, bmax.y + pad, bmax.z + pad};
    const float o[3] = {ray.origin.x, ray.origin.y, ray.origin.z}, k[3] = {inv.x, inv.y, inv.z};
    float t1 = (lo[0] - o[0]) * k[0], t2 = (hi[0] - o[0]) * k[0];
    float tmin = fminf(t1, t2), tmax = fmaxf(t1, t2);
    for (int a = 1; a < 3; ++a) {
        t1 = (lo[a] - o[a]) * k[a];
        t2 = (hi[a] - o[a]) * k[a];
        tmin = fmaxf(tmin, fminf(t1, t2));
        tmax = fminf(tmax, fmaxf(t1, t2));
    }
    if (tmax < 0.0f || tmin > tMax || tmin > tmax) return false;
    if (tNear) *tNear = tmin;
    return true;
}

// ---- Moller-Trumbore: src/mesh_intersection.cuh:15-46 ------------------------------------------------
inline PreliminaryHitInfo intersectTriangleIndexed(const Ray& ray, const Vec3& v0, const Vec3& v1, const Vec3& v2) {
    PreliminaryHitInfo miss;
    const Vec3 e1 = v1 - v0, e2 = v2 - v0;
    const Vec3 pv = cross(ray.direction, e2);
    const float det = dot(e1, pv);
    if (fabsf(det) < 1e-8f) return miss;
    const float inv = 1.0f / det;
    const Vec3 tv = ray.origin - v0;
    const float u = dot(tv, pv) * inv;
    if (u < 0.0f || u > 1.0f) return miss;
    const Vec3 qv = cross(tv, e1);
    const float v = dot(ray.direction, qv) * inv;
    if (v < 0.0f || u + v > 1.0f) return miss;
    const float t = dot(e2, qv) * inv;
    if (t <= 1e-8f) return miss;
    PreliminaryHitInfo h;
    h.t = t;
    h.u = u;
    h.v = v;
    return h;
}

// ---- textures and materials: src/mesh_intersection.cuh:53-152, 244-337 ------------------------------
inline Vec3 sampleTextureRawDev(const TextureDeviceView& tex, float u, float v, bool nearest) {
    if (!tex.pixels || tex.width <= 0 || tex.height <= 0 || tex.channels <= 0) return Vec3(-1.0f, -1.0f, -1.0f);
    u -= floorf(u);
    v -= floorf(v);
    auto at = [&](int xi, int yi) {
        const unsigned char* p = tex.pixels + (yi * tex.width + xi) * tex.channels;
        return Vec3(p[0] * (1.0f / 255.0f), tex.channels > 1 ? p[1] * (1.0f / 255.0f) : 0.0f,
                    tex.channels > 2 ? p[2] * (1.0f / 255.0f) : 0.0f);
    };
    auto clampi = [](int a, int lo, int hi) { return a < lo ? lo : (a > hi ? hi : a); };
    if (nearest)
        return at(clampi(static_cast<int>(u * static_cast<float>(tex.width)), 0, tex.width - 1),
                  clampi(static_cast<int>(v * static_cast<float>(tex.height)), 0, tex.height - 1));
    const float x = u * static_cast<float>(tex.width - 1), y = v * static_cast<float>(tex.height - 1);
    const int x0 = static_cast<int>(floorf(x)), y0 = static_cast<int>(floorf(y));
    const int x1 = x0 + 1 < tex.width - 1 ? x0 + 1 : tex.width - 1, y1 = y0 + 1 < tex.height - 1 ? y0 + 1 : tex.height - 1;
    const float tx = x - static_cast<float>(x0), ty = y - static_cast<float>(y0);
    return lerp(lerp(at(x0, y0), at(x1, y0), tx), lerp(at(x0, y1), at(x1, y1), tx), ty);
}

inline int getMaterialIdForTriangle(uint32_t tri, const MeshDeviceView& mesh) {
    if (!mesh.hasMeshMaterials || mesh.numPrimitives <= 0) return -1;
    uint32_t prim = 0;
    if (mesh.materialMap) {  // last primitive whose first triangle is <= tri (:53-70)
        uint32_t lo = 0, hi = static_cast<uint32_t>(mesh.numPrimitives);
        while (lo < hi) {
            const uint32_t mid = lo + (hi - lo) / 2;
            if (mesh.materialMap[mid] <= tri) lo = mid + 1;
            else hi = mid;
        }
        prim = lo > 0 ? lo - 1 : 0;
    }
    if (prim < static_cast<uint32_t>(mesh.numPrimitives) && mesh.materialIds) return mesh.materialIds[prim];
    return -1;
}

struct ResolvedMaterial {
    Vec3 base_color;
    float metallic = 0.0f, roughness = 0.0f, specular = 0.0f;
};
inline float sampleMaterialParam(const MaterialParam& p, Vec2 uv, const MeshDeviceView& mesh) {
    if (!p.textured || p.textureId >= static_cast<uint32_t>(mesh.numTextures) || !mesh.textures) return p.value;
    const Vec3 s = sampleTextureRawDev(mesh.textures[p.textureId], uv.x, uv.y, mesh.textureNearest != 0);
    if (s.x < 0.0f) return p.value;
    return p.channel == 1 ? s.y : (p.channel == 2 ? s.z : s.x);
}
inline ResolvedMaterial resolveMaterial(const Material& mat, Vec2 uv, const MeshDeviceView& mesh) {
    ResolvedMaterial r;  // only the fields the kernels consume; the sRGB decode is commented out upstream (:288-290)
    r.base_color = mat.base_color.value;
    if (mat.base_color.textured && mat.base_color.textureId < static_cast<uint32_t>(mesh.numTextures) && mesh.textures) {
        const Vec3 raw = sampleTextureRawDev(mesh.textures[mat.base_color.textureId], uv.x, uv.y, mesh.textureNearest != 0);
        if (raw.x >= 0.0f) r.base_color = raw;
    }
    r.metallic = sampleMaterialParam(mat.metallic, uv, mesh);
    r.roughness = sampleMaterialParam(mat.roughness, uv, mesh);
    r.specular = sampleMaterialParam(mat.specular, uv, mesh);
    return r;
}

// ---- hit record: src/mesh_intersection.cuh:91-103, 168-238 -------------------------------------------
inline HitInfo computeHitData(const PreliminaryHitInfo& pi, uint32_t tri, const Ray&, const MeshDeviceView& mesh) {
    HitInfo h;
    h.t = pi.t;
    const uint3 id = mesh.indices[tri];
    const Vec3 a = mesh.vertices[id.x], b = mesh.vertices[id.y], c = mesh.vertices[id.z];
    const float w = 1.0f - pi.u - pi.v;
    h.geometricNormal = normalize(cross(b - a, c - a));
    h.shadingNormal = (mesh.hasNormals && mesh.normals)
                          ? normalize(mesh.normals[id.x] * w + mesh.normals[id.y] * pi.u + mesh.normals[id.z] * pi.v)
                          : h.geometricNormal;
    h.position = a * w + b * pi.u + c * pi.v;
    h.uv = (mesh.hasTexcoords && mesh.texcoords)
               ? mesh.texcoords[id.x] * w + mesh.texcoords[id.y] * pi.u + mesh.texcoords[id.z] * pi.v
               : Vec2(pi.u, pi.v);
    h.materialId = getMaterialIdForTriangle(tri, mesh);
    if (h.materialId >= 0 && h.materialId < mesh.numMaterials && mesh.materials) {
        const Material& m = mesh.materials[h.materialId];
        if (m.normal.textured && m.normal.textureId < static_cast<uint32_t>(mesh.numTextures)) {
            const Vec3 tn = sampleTextureRawDev(mesh.textures[m.normal.textureId], h.uv.x, h.uv.y, mesh.textureNearest != 0);
            if (tn.x >= 0.0f) {
                const Vec3 ts = normalize(tn * 2.0f - Vec3(1.0f, 1.0f, 1.0f));
                Vec3 bt;
                if (fabsf(h.shadingNormal.x) < 0.6f) bt.x = 1.0f;
                else if (fabsf(h.shadingNormal.y) < 0.6f) bt.y = 1.0f;
                else bt.z = 1.0f;
                const Vec3 tg = normalize(cross(bt, h.shadingNormal));
                bt = normalize(cross(h.shadingNormal, tg));
                h.shadingNormal = normalize(tg * ts.x + bt * ts.y + h.shadingNormal * ts.z);
            }
        }
    }
    return h;
}

// ---- closest-hit DFS: src/mesh_traversal.cuh:107-231 --------------------------------------------------
enum class TraceMode { ANY, FORWARD_ONLY, BACKWARD_ONLY };

inline bool traceMeshWithMode(const Ray& ray, MeshDeviceView mesh, HitInfo* out, TraceMode mode, bool, const Material&) {
    if (mesh.numBvhNodes <= 0 || mesh.numTriangles <= 0) return false;
    PreliminaryHitInfo best;
    uint32_t bestTri = 0;
    const Vec3 inv(1.0f / ray.direction.x, 1.0f / ray.direction.y, 1.0f / ray.direction.z);
    constexpr int kDepth = 256;
    int todo[kDepth];
    int top = 0;
    todo[top++] = 0;
    auto inRange = [&](int n) { return n >= 0 && n < mesh.numBvhNodes; };
    auto push = [&](int n) {
        if (top < kDepth) todo[top++] = n;  // overflow silently drops, as upstream (:202-207)
    };
    while (top > 0) {
        const int ni = todo[--top];
        if (!inRange(ni)) continue;
        const BvhNode node = mesh.bvhNodes[ni];
        float tn = 0.0f;
        if (!intersectAabb(ray, inv, node.boundsMin, node.boundsMax, best.t, &tn)) continue;  // re-test vs current best
        if (node.isLeaf) {
            for (int i = node.first; i < node.first + node.count; ++i) {
                const uint3 id = mesh.indices[i];
                const Vec3 a = mesh.vertices[id.x], b = mesh.vertices[id.y], c = mesh.vertices[id.z];
                const float facing = dot(normalize(cross(b - a, c - a)), ray.direction);
                if (mode == TraceMode::FORWARD_ONLY && facing >= 0.0f) continue;
                if (mode == TraceMode::BACKWARD_ONLY && facing <= 0.0f) continue;
                const PreliminaryHitInfo pi = intersectTriangleIndexed(ray, a, b, c);
                if (pi.valid() && pi.t > 1e-6f && pi.t < best.t) {
                    best = pi;
                    bestTri = static_cast<uint32_t>(i);
                }
            }
            continue;
        }
        float nearL = 0.0f, nearR = 0.0f;
        const bool hitL = inRange(node.left) && intersectAabb(ray, inv, mesh.bvhNodes[node.left].boundsMin,
                                                              mesh.bvhNodes[node.left].boundsMax, best.t, &nearL);
        const bool hitR = inRange(node.right) && intersectAabb(ray, inv, mesh.bvhNodes[node.right].boundsMin,
                                                               mesh.bvhNodes[node.right].boundsMax, best.t, &nearR);
        if (hitL && hitR) {  // far child first so the near one is popped next (:195-203)
            const bool rightFirst = nearR < nearL;
            push(rightFirst ? node.left : node.right);
            push(rightFirst ? node.right : node.left);
        } else if (hitL) {
            push(node.left);
        } else if (hitR) {
            push(node.right);
        }
    }
    if (!best.valid()) return false;
    best.primIdx = bestTri;
    if (out) *out = computeHitData(best, bestTri, ray, mesh);
    return true;
}

inline bool traceMesh(const Ray& ray, MeshDeviceView mesh, HitInfo* out, bool cullBackfaces, const Material& g) {
    return traceMeshWithMode(ray, mesh, out, cullBackfaces ? TraceMode::FORWARD_ONLY : TraceMode::ANY, false, g);
}
