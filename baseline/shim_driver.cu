// baseline/shim_driver.cu - RUNS the source-compatible C++ class of the boundary (include/nrb200_renderer.hpp, the
// drop-in for src/cuda_renderer_neural.h:84-124) over libnrb200.so, against the reference's OWN unmodified
// scene.h / mesh.h / renderer.h / config_loader.h and its own Mesh::buildBvh (src/mesh_bvh.cpp), in evaluate's call
// sequence (src/evaluate.cu:211-296: construct, setConstantNeuralColor, resize, setBounceCount, setLambertView,
// setEnvmapRotation, setUseNeuralQuery, loadWeightsFromFile, setCameraBasis, then per pass setUseNeuralQuery /
// setClassicMeshIndex / resetSamples / setSamplesPerPixel / render ... and a blocking cudaMemcpy from devicePixels()).
// The two private host arrays the class needs are reached through `#define private public` here
// instead of the two accessor patches.  Built by baseline/Makefile (needs /root/reference); test infrastructure.
//
//   shim_driver <scene.pack> <out_dir>
#define PACK_DEFINE_EXR_STUBS
#include "pack_reader.h"

#define NRB_MESH_BVH_NODES(mesh) ((mesh).bvhNodes_)
#define NRB_ENV_PIXELS(env) ((env).pixels_)
#define NRB_ENV_WIDTH(env) ((env).width_)
#define NRB_ENV_HEIGHT(env) ((env).height_)
#define NRB_ENV_STRENGTH(env) ((env).strength_)
#define PROFILE_KERNELS
#include "nrb200_renderer.hpp"

using namespace pack;

int main(int argc, char** argv) {
    if (argc < 3) {
        std::fprintf(stderr, "usage: shim_driver <scene.pack> <out_dir>\n");
        return 2;
    }
    const std::string outDir = argv[2];
    Reader r(argv[1]);
    const Params p = openPack(r);
    Scene scene;
    Mesh* meshes[4] = {&scene.originalMesh(), &scene.innerShell(), &scene.outerShell(), &scene.additionalMesh()};
    std::vector<char> weights;
    loadScene(r, p, scene, weights);
    for (int i = 0; i < 4; ++i)
        if (meshes[i]->numTriangles() > 0) dumpBvh(outDir + "/bvh_" + std::to_string(i) + ".bin", *meshes[i]);

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
    renderer.setSwapParamOrder(p.swapOrder != 0);
    renderer.setUseNeuralQuery(p.neural != 0);
    if (!weights.empty()) {
        const std::string wpath = outDir + "/_weights.bin";
        dump(wpath, weights.data(), weights.size());
        if (!renderer.loadWeightsFromFile(wpath)) return 3;
        if (renderer.loadWeightsFromFile(outDir + "/_no_such_file.bin")) return 4;  // must report false, not exit
    }
    RenderBasis basis;
    basis.forward = Vec3(p.fwd[0], p.fwd[1], p.fwd[2]);
    basis.right = Vec3(p.right[0], p.right[1], p.right[2]);
    basis.up = Vec3(p.up[0], p.up[1], p.up[2]);
    basis.fovY = p.fovY;
    renderer.setCameraBasis(basis);
    const Vec3 camPos(p.camPos[0], p.camPos[1], p.camPos[2]);
    const size_t px = static_cast<size_t>(p.width) * p.height;
    std::vector<uchar4> frame(px);

    // evaluate's two passes: ground truth first, then neural (when the pack asks for it), each accumFrames batches
    for (int pass = 0; pass < (p.neural ? 2 : 1); ++pass) {
        renderer.setUseNeuralQuery(pass == 1);
        renderer.setClassicMeshIndex(p.classicMeshIndex);
        renderer.resetSamples();
        for (int i = 0; i < (p.accumFrames > 0 ? p.accumFrames : 1); ++i) {
            renderer.setSamplesPerPixel(p.spp);
            renderer.render(camPos);
        }
        cudaMemcpy(frame.data(), renderer.devicePixels(), px * sizeof(uchar4), cudaMemcpyDeviceToHost);  // evaluate.cu:266
        dump(outDir + (pass == 0 ? "/shim_gt.rgba" : "/shim_neural.rgba"), frame.data(), px * 4);
        const KernelTimings& t = renderer.lastFrameTimings();
        std::printf("pass %d: rays %d neural calls %llu total %.3f ms", pass, t.rayCount,
                    static_cast<unsigned long long>(t.neuralRayCalls), t.totalMs());
        for (int k = 0; k < KID_COUNT; ++k)
            if (t.ms[k] > 0.0) std::printf(" | %s %.3f", KernelTimings::name(k), t.ms[k]);
        std::printf("\n");
    }
    std::printf("shim_driver: %dx%d spp %d bounces %d params %zu bytes ok\n", renderer.width(), renderer.height(),
                renderer.samplesPerPixel(), renderer.bounceCount(), renderer.paramsBytes());
    return 0;
}
