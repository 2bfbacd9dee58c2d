// nrb200_renderer.hpp — source-compatible replacement of the reference's `class RendererNeural`
// (src/cuda_renderer_neural.h:84-124) implemented over the C ABI of libnrb200.so (nrb200.h).
//
// A maintainer of the reference drops this header in place of cuda_renderer_neural.h (and removes
// cuda_renderer_neural.cu + tiny-cuda-nn from the build); evaluate.cu / viewer.cu compile unchanged.
// It is compiled against the reference's OWN scene.h / mesh.h / renderer.h / config_loader.h, which
// stay as they are, except that two private host arrays must be readable:
//     Mesh::bvhNodes_            -> add  const std::vector<BvhNode>& bvhNodes() const { return bvhNodes_; }
//     EnvironmentMap::pixels_    -> add  hostPixels(), width(), height(), rotation(), strength()
// The access expressions are macros so that a tree that exposes them differently can override them.
//
// Behaviour mirrored from the reference: CUDA failures print and exit(1) (checkCuda,
// src/cuda_renderer_neural.cu:1619-1624); loadWeightsFromFile returns false with a message on
// open/size mismatch (:1752-1789); render() silently returns when the size is not set (:2086-2091).
#pragma once

#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include <cuda_runtime.h>  // uchar4

#include "config_loader.h"  // NeuralNetworkConfig   (reference header)
#include "renderer.h"       // RenderBasis            (reference header)
#include "scene.h"          // Scene, Mesh, Material  (reference header)

#include "nrb200.h"

#ifndef NRB_MESH_BVH_NODES
#define NRB_MESH_BVH_NODES(mesh) ((mesh).bvhNodes())
#endif
#ifndef NRB_ENV_PIXELS
#define NRB_ENV_PIXELS(env) ((env).hostPixels())
#define NRB_ENV_WIDTH(env) ((env).width())
#define NRB_ENV_HEIGHT(env) ((env).height())
#define NRB_ENV_STRENGTH(env) ((env).strength())
#endif

// Per-kernel GPU timing, as in the reference (src/cuda_renderer_neural.h:17-66): only with -DPROFILE_KERNELS.
// Same enumerators, names and struct; the values come from nrb_get_timings (NRB_OPT_PROFILE event pairs), with this
// library's fused kernels attributed to the bucket of their dominant part (csrc/nrb_api.cu:timing_bucket).
#ifdef PROFILE_KERNELS
enum KernelId {
    KID_primaryTrace, KID_bounceTrace, KID_shellTrace, KID_neuralForward, KID_additionalMeshPrimary,
    KID_selectClosestPrimary, KID_initPathState, KID_sampleBounce, KID_earlyTermination, KID_additionalMeshBounce,
    KID_selectClosestBounce, KID_integrateBounce, KID_finalize, KID_COUNT
};
static_assert(KID_COUNT == NRB_TIMING_BUCKETS, "KernelId buckets");

struct KernelTimings {
    double ms[KID_COUNT] = {};
    int rayCount = 0;             // width * height * samplesPerPixel
    uint64_t neuralRayCalls = 0;  // sum of active-ray counts across all neural forward passes

    static const char* name(int kid) {
        static const char* kNames[KID_COUNT] = {
            "GT primary intersect", "GT bounce intersect", "Shell intersection", "Neural forward pass",
            "Additional mesh (primary)", "Select closest (primary)", "Init path state", "Sample bounce dirs",
            "Bounce early term.", "Additional mesh (bounce)", "Select closest (bounce)", "Integrate bounce",
            "Finalize / lambert",
        };
        return (kid >= 0 && kid < KID_COUNT) ? kNames[kid] : "?";
    }
    double totalMs() const {
        double t = 0.0;
        for (int i = 0; i < KID_COUNT; ++i) t += ms[i];
        return t;
    }
};
#endif  // PROFILE_KERNELS

class RendererNeural final {
 public:
    explicit RendererNeural(Scene& scene, const NeuralNetworkConfig* nnConfig = nullptr) : scene_(&scene) {
        nrb_nn_config cfg{};  // zeros select the reference's hard-coded network shape
        if (nnConfig) {
            cfg.log2_hashmap_size = nnConfig->log2_hashmap_size;
            cfg.base_resolution = nnConfig->base_resolution;
        }
        check(nrb_create(&cfg, -1, &r_), "nrb_create");
#ifdef PROFILE_KERNELS
        nrb_set_option(r_, NRB_OPT_PROFILE, 1);
#endif
    }
    ~RendererNeural() { nrb_destroy(r_); }
    RendererNeural(const RendererNeural&) = delete;
    RendererNeural& operator=(const RendererNeural&) = delete;

    void resize(int width, int height) { check(nrb_resize(r_, width, height), "nrb_resize"); }
    void setCameraBasis(const RenderBasis& b) {
        check(nrb_set_camera_basis(r_, &b.forward.x, &b.right.x, &b.up.x, b.fovY), "nrb_set_camera_basis");
    }
    void render(const Vec3& camPos) {
        syncScene();
        check(nrb_render(r_, &camPos.x), "nrb_render");
    }
    uchar4* devicePixels() const { return static_cast<uchar4*>(nrb_device_pixels(r_)); }

    void setUseNeuralQuery(bool enabled) { useNeuralQuery_ = enabled; nrb_set_option(r_, NRB_OPT_USE_NEURAL_QUERY, enabled); }
    bool useNeuralQuery() const { return useNeuralQuery_; }
    void setUseHardwareRT(bool v) { useHardwareRT_ = v; nrb_set_option(r_, NRB_OPT_USE_HARDWARE_RT, v); }
    bool useHardwareRT() const { return useHardwareRT_; }
    bool loadWeightsFromFile(const std::string& path) {
        if (nrb_load_weights_file(r_, path.c_str(), swapParamOrder_ ? 1 : 0) != NRB_OK) {
            std::fprintf(stderr, "%s\n", nrb_last_error());
            return false;
        }
        return true;
    }
    void setSwapParamOrder(bool v) { swapParamOrder_ = v; }
    bool swapParamOrder() const { return swapParamOrder_; }

    void setSamplesPerPixel(int samples) { samplesPerPixel_ = samples; nrb_set_option(r_, NRB_OPT_SAMPLES_PER_PIXEL, samples); }
    void setBounceCount(int count) { bounceCount_ = count; nrb_set_option(r_, NRB_OPT_BOUNCE_COUNT, count); }
    void setLambertView(bool enabled) { nrb_set_option(r_, NRB_OPT_LAMBERT_VIEW, enabled); }
    void setClassicMeshIndex(int index) { classicMeshIndex_ = index; nrb_set_option(r_, NRB_OPT_CLASSIC_MESH_INDEX, index); }
    int classicMeshIndex() const { return classicMeshIndex_; }
    void setEnvmapRotation(float degrees) { envmapRotation_ = degrees; nrb_set_envmap_rotation(r_, degrees); }
    float envmapRotation() const { return envmapRotation_; }
    void setConstantNeuralColor(bool use, Vec3 color) {
        nrb_set_option(r_, NRB_OPT_USE_CONSTANT_NEURAL_COLOR, use);
        nrb_set_constant_neural_color(r_, &color.x);
    }
    void setDirectEnvColor(bool use, Vec3 color) {
        nrb_set_option(r_, NRB_OPT_USE_DIRECT_ENV_COLOR, use);
        nrb_set_direct_env_color(r_, &color.x);
    }
    void setUseAdditionalMesh(bool enabled) { nrb_set_option(r_, NRB_OPT_USE_ADDITIONAL_MESH, enabled); }
    void resetSamples() { nrb_reset_samples(r_); }

    int samplesPerPixel() const { return samplesPerPixel_; }
    int bounceCount() const { return bounceCount_; }
    size_t paramsBytes() const { return nrb_params_bytes(r_); }
    int width() const { return nrb_width(r_); }
    int height() const { return nrb_height(r_); }

#ifdef PROFILE_KERNELS
    // The reference fills lastFrameTimings_ at the end of render() after a device synchronise (:2738-2751); here the
    // events are read on demand.
    const KernelTimings& lastFrameTimings() const {
        uint64_t calls = 0;
        check(nrb_get_timings(r_, lastFrameTimings_.ms, &lastFrameTimings_.rayCount, &calls), "nrb_get_timings");
        lastFrameTimings_.neuralRayCalls = calls;
        return lastFrameTimings_;
    }
#endif

    // The reference re-uploads a mesh when one of its private dirty flags is set (src/mesh.cu:91-222).  The flags are
    // not readable from outside, so a slot is uploaded again when what can be observed of it changed - triangle,
    // vertex and node counts or the address of a host array (a reload, clear(), or rebuilt BVH moves them) - and
    // callers that edit vertex data in place call markSceneDirty().
    void markSceneDirty() { sceneUploaded_ = false; for (auto& s : slotSeen_) s = SlotSeen{}; }
    nrb_renderer* handle() const { return r_; }

 private:
    static void check(int rc, const char* what) {
        if (rc != NRB_OK) {
            std::fprintf(stderr, "CUDA error (%s): %s\n", what, nrb_last_error());
            std::exit(1);
        }
    }

    struct SlotSeen {
        const void* vertices = nullptr;
        const void* indices = nullptr;
        int triangles = -1, numVertices = -1, nodes = -1, materials = -1, textures = -1;
        bool operator==(const SlotSeen& o) const {
            return vertices == o.vertices && indices == o.indices && triangles == o.triangles && numVertices == o.numVertices &&
                   nodes == o.nodes && materials == o.materials && textures == o.textures;
        }
    };
    static SlotSeen observe(const Mesh& m) {
        SlotSeen s;
        s.vertices = m.vertices_.data(); s.indices = m.indices_.data();
        s.triangles = m.numTriangles(); s.numVertices = m.numVertices(); s.nodes = m.nodeCount();
        s.materials = static_cast<int>(m.materials_.size()); s.textures = m.textureCount();
        return s;
    }

    void uploadMesh(int slot, Mesh& mesh) {
        if (mesh.numTriangles() > 0 && !mesh.hasBvh()) mesh.buildBvh();  // CPU SAH build of the reference; defines primitive ids
        const SlotSeen now = observe(mesh);
        if (sceneUploaded_ && now == slotSeen_[slot]) return;
        slotSeen_[slot] = now;
        if (mesh.numTriangles() == 0) {
            check(nrb_update_mesh(r_, slot, nullptr), "nrb_update_mesh");
            return;
        }
        const auto& nodes = NRB_MESH_BVH_NODES(mesh);
        static_assert(sizeof(BvhNode) == sizeof(nrb_bvh_node), "BvhNode layout");
        static_assert(sizeof(Material) == sizeof(nrb_material), "Material layout");
        nrb_mesh_desc d{};
        d.vertices = &mesh.vertices_[0].x;
        d.normals = mesh.normals_.empty() ? nullptr : &mesh.normals_[0].x;
        d.texcoords = mesh.texcoords_.empty() ? nullptr : &mesh.texcoords_[0].x;
        d.indices = &mesh.indices_[0].x;
        d.num_triangles = mesh.numTriangles();
        d.num_vertices = mesh.numVertices();
        d.nodes = reinterpret_cast<const nrb_bvh_node*>(nodes.data());
        d.num_nodes = static_cast<int>(nodes.size());
        d.materials = mesh.materials_.empty() ? nullptr : reinterpret_cast<const nrb_material*>(mesh.materials_.data());
        d.num_materials = static_cast<int>(mesh.materials_.size());
        d.material_map = mesh.materialMap_.empty() ? nullptr : mesh.materialMap_.data();
        d.material_ids = mesh.materialIds_.empty() ? nullptr : mesh.materialIds_.data();
        d.num_primitives = static_cast<int>(mesh.materialMap_.size());
        std::vector<nrb_texture> tex(mesh.textures_.size());
        for (size_t i = 0; i < tex.size(); ++i) {
            const MeshTexture& t = mesh.textures_[i];
            tex[i] = nrb_texture{t.pixels.data(), t.width, t.height, t.channels, t.colorSpace == ColorSpace::SRGB ? 1 : 0};
        }
        d.textures = tex.empty() ? nullptr : tex.data();
        d.num_textures = static_cast<int>(tex.size());
        d.texture_nearest = mesh.textureNearest() ? 1 : 0;
        d.has_mesh_materials = mesh.hasMeshMaterials() ? 1 : 0;
        const Vec3 mn = mesh.boundsMin(), mx = mesh.boundsMax();
        d.bounds_min[0] = mn.x; d.bounds_min[1] = mn.y; d.bounds_min[2] = mn.z;
        d.bounds_max[0] = mx.x; d.bounds_max[1] = mx.y; d.bounds_max[2] = mx.z;
        check(nrb_update_mesh(r_, slot, &d), "nrb_update_mesh");
    }

    void syncScene() {
        // global material can be edited between frames by the callers (evaluate.cu:181-209)
        check(nrb_set_global_material(r_, reinterpret_cast<const nrb_material*>(&scene_->globalMaterial())),
              "nrb_set_global_material");
        uploadMesh(NRB_MESH_ORIGINAL, scene_->originalMesh());
        uploadMesh(NRB_MESH_INNER_SHELL, scene_->innerShell());
        uploadMesh(NRB_MESH_OUTER_SHELL, scene_->outerShell());
        uploadMesh(NRB_MESH_ADDITIONAL, scene_->additionalMesh());
        if (sceneUploaded_) return;
        EnvironmentMap& env = scene_->environment();
        if (env.isValid()) {
            check(nrb_update_env(r_, &NRB_ENV_PIXELS(env)[0].x, NRB_ENV_WIDTH(env), NRB_ENV_HEIGHT(env), NRB_ENV_STRENGTH(env)),
                  "nrb_update_env");
        } else {
            check(nrb_update_env(r_, nullptr, 0, 0, 1.0f), "nrb_update_env");
        }
        sceneUploaded_ = true;
    }

    Scene* scene_ = nullptr;
    nrb_renderer* r_ = nullptr;
    bool sceneUploaded_ = false;
    SlotSeen slotSeen_[NRB_MESH_SLOT_COUNT];
#ifdef PROFILE_KERNELS
    mutable KernelTimings lastFrameTimings_;
#endif
    bool useNeuralQuery_ = false, useHardwareRT_ = false, swapParamOrder_ = true;
    int samplesPerPixel_ = 1, bounceCount_ = 0, classicMeshIndex_ = 0;
    float envmapRotation_ = 0.0f;
};
