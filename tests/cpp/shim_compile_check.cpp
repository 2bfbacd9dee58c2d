// Compile check of include/nrb200_renderer.hpp against the reference's UNMODIFIED headers (used by
// tests/test_host_cpu.py where /root/reference exists).  The two private host arrays the shim needs
// are reached here with the `#define private public` harness trick instead of the one-line accessors
// INTEGRATION.md asks a maintainer to add.
#include <cuda_runtime.h>
#define private public
#include "scene.h"
#undef private
#define NRB_MESH_BVH_NODES(mesh) ((mesh).bvhNodes_)
#define NRB_ENV_PIXELS(env) ((env).pixels_)
#define NRB_ENV_WIDTH(env) ((env).width_)
#define NRB_ENV_HEIGHT(env) ((env).height_)
#define NRB_ENV_STRENGTH(env) ((env).strength_)
#include "nrb200_renderer.hpp"

// The call sequence of the reference's evaluate driver (src/evaluate.cu:211-296).
int drive(Scene& scene, const NeuralNetworkConfig& nn, const RenderBasis& basis, const Vec3& cam) {
    RendererNeural renderer(scene, &nn);
    renderer.resize(1920, 1080);
    renderer.setBounceCount(3);
    renderer.setEnvmapRotation(0.0f);
    if (!renderer.loadWeightsFromFile("weights.bin")) return 1;
    renderer.setCameraBasis(basis);
    renderer.setUseNeuralQuery(false);
    renderer.setSamplesPerPixel(8);
    renderer.render(cam);
    renderer.setUseNeuralQuery(true);
    renderer.render(cam);
    return renderer.devicePixels() != nullptr && renderer.paramsBytes() > 0 ? 0 : 2;
}
