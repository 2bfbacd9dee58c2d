// nrb_scene.cuh — device-side scene layout of libnrb200 (what nrb_update_mesh builds).
//
// The reference hands kernels a MeshDeviceView of 44-byte AoS BvhNode records and an index
// indirection per triangle (src/mesh.h:45-78, src/bvh_data.h:5-13).  Here every mesh is re-laid
// out once at upload into
//   * NodePair[]  : one 64-byte, 16-byte-aligned record per INNER node holding both children's
//                   boxes and references -> 4 x LDG.128 per traversal step (the reference reads
//                   2 x 44 B as 22 scalar loads and re-reads each box when the child is popped);
//   * TriBlock[]  : one 48-byte record per triangle in BVH leaf order (= primitive id order):
//                   v0, e1 = v1-v0, e2 = v2-v0 and the unit geometric normal, pre-computed with
//                   the reference's own operation order -> 3 x LDG.128 per candidate triangle
//                   instead of 1 index + 3 vertex gathers + a normalize;
// plus the reference's attribute arrays (used once per ray for the hit record).
#pragma once

#include "nrb_math.cuh"

namespace nrb {

// Child reference: >= 0 -> index into NodePair[]; < 0 -> leaf, first triangle = ~ref.
struct __align__(16) NodePair {
    // (x, y) and (z, z) couples: every 8-byte half of a record meets the same couple of ray constants, so
    // the 12 subtractions and 12 multiplications of the two slab tests are 6 + 6 packed fp32x2 instructions
    float4 a;  // lmin.x lmin.y lmax.x lmax.y
    float4 b;  // rmin.x rmin.y rmax.x rmax.y
    float4 c;  // lmin.z lmax.z rmin.z rmax.z
    int4 d;    // leftRef rightRef leftCone rightCone
};

// Normal cone of a child's subtree, one 32-bit word: axis as three snorm8 (x | y << 8 | z << 16, value / 127) and a
// threshold s as unorm8 (<< 24, value / 255; 255 = no cone).  Built at upload (nrb_api.cu:relayout_mesh) such that
//   dot(axis, d) >=  s * |d|  implies  n . d > 0 for EVERY triangle normal n of the subtree (all back-facing: a
//                                       FORWARD_ONLY query rejects each of them, src/mesh_traversal.cuh:166-169),
//   dot(axis, d) <= -s * |d|  implies  n . d < 0 for every one (all front-facing: BACKWARD_ONLY rejects them all),
// with margins that cover the 8-bit quantisation and fp32 rounding of both dot products.  Skipping such a subtree for
// that mode removes only triangles its facing test would have skipped, so results are unchanged; what it buys is
// that a FORWARD query along a ray that LEAVES a closed shell stops visiting everything the ray pierces.
constexpr uint32_t kNoCone = 0xff000000u;
__host__ __device__ __forceinline__ float cone_dot(uint32_t cone, float dx, float dy, float dz) {
    const float k = 1.0f / 127.0f;
    const float ax = static_cast<float>(static_cast<signed char>(cone & 0xffu)) * k;
    const float ay = static_cast<float>(static_cast<signed char>((cone >> 8) & 0xffu)) * k;
    const float az = static_cast<float>(static_cast<signed char>((cone >> 16) & 0xffu)) * k;
    return ax * dx + ay * dy + az * dz;
}
__host__ __device__ __forceinline__ float cone_threshold(uint32_t cone) { return static_cast<float>(cone >> 24) * (1.0f / 255.0f); }
// bit 0: no triangle of the subtree can pass a FORWARD_ONLY facing test along d; bit 1: none can pass BACKWARD_ONLY
__host__ __device__ __forceinline__ int cone_culls(uint32_t cone, float dx, float dy, float dz, float dlen) {
    if ((cone >> 24) == 0xffu) return 0;
    const float c = cone_dot(cone, dx, dy, dz), s = cone_threshold(cone) * dlen;
    return (c >= s ? 1 : 0) | (c <= -s ? 2 : 0);
}
static_assert(sizeof(NodePair) == 64, "NodePair");

struct __align__(16) TriBlock {
    float4 p;  // v0.xyz, n.x
    float4 q;  // e1.xyz, n.y
    float4 r;  // e2.xyz, n.z
};
static_assert(sizeof(TriBlock) == 48, "TriBlock");

// = Material (reference src/material.h:9-117), same 180-byte layout as nrb_material.
struct MatParam {
    float value;
    uint32_t tex;
    uint8_t channel, textured, pad_[2];
};
struct MatParam3 {
    float value[3];
    uint32_t tex;
    uint8_t textured, pad_[3];
};
struct MaterialDev {
    MatParam3 base_color;
    MatParam metallic, roughness, specular, specular_tint, anisotropy, sheen, sheen_tint, clearcoat, clearcoat_gloss;
    MatParam3 normal, base_emission;
    float emission_scale, ior, specular_transmission;
};
static_assert(sizeof(MaterialDev) == 180, "Material layout");

struct TexDev {
    const unsigned char* pixels;
    int width, height, channels, srgb;
};

struct MeshDev {
    const NodePair* nodes;
    const TriBlock* tris;
    const float* vertices;    // 3 per vertex (reference layout)
    const float* normals;     // or null
    const float* texcoords;   // or null
    const uint32_t* indices;  // 3 per triangle, leaf order
    const MaterialDev* materials;
    const uint32_t* materialMap;
    const int* materialIds;
    const TexDev* textures;
    int numTriangles, numPairs;
    int rootRef, rootCount;   // root as a child reference (a single-leaf mesh has rootRef < 0)
    float rootMin[3], rootMax[3];
    int numPrimitives, numMaterials, numTextures, textureNearest, hasMeshMaterials;
};

struct EnvDev {
    const float* pixels;  // fp32 RGB triplets (reference EnvironmentDeviceView, src/scene.h:10-16)
    int width, height;
    float rotation, strength;
};

// The constants of one frame (the reference passes a 332-byte RenderParams by value to every
// kernel, src/render_params.h; this is the same information with the material resolved).
struct FrameParams {
    V3 camPos, camForward, camRight, camUp;
    V3 outerMin, outerInvExtent;
    V3 constantNeuralColor, directEnvColor;
    float tanHalfFov, aspect;   // tanf(fovY/2) evaluated on the host, as the oracle does
    float maxRadiance, sceneScale;
    int useConstantNeuralColor, useDirectEnvColor;
    int maxBounces, width, height, pixelCount, spp;
    uint32_t sampleOffset;
    // tile ownership (multi-GPU): local row r -> global row
    int bandRows, rank, world, localHeight;
    int localPixels;            // width * localHeight
    int nElems;                 // localPixels * spp
    int nPoints, inputDims;
};

enum { TRACE_ANY = 0, TRACE_FORWARD = 1, TRACE_BACKWARD = 2 };
enum { NRB_SLOT_ORIGINAL = 0, NRB_SLOT_INNER = 1, NRB_SLOT_OUTER = 2, NRB_SLOT_ADDITIONAL = 3 };

struct Hit {
    float t, u, v;
    uint32_t prim;
};

}  // namespace nrb
