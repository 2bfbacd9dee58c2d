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
#define PACK_DEFINE_EXR_STUBS
#include "pack_reader.h"
#define private public
#include "cuda_renderer_neural.h"
#undef private

using namespace pack;

int main(int argc, char** argv) {
    if (argc < 3) {
        std::fprintf(stderr, "usage: ref_render <scene.pack> <out_dir> [warmup timed]\n");
        return 2;
    }
    const std::string outDir = argv[2];
    const int warmup = argc > 3 ? std::atoi(argv[3]) : 0, timed = argc > 4 ? std::atoi(argv[4]) : 0;
    Reader r(argv[1]);
    const Params p = openPack(r);
    Scene scene;
    Mesh* meshes[4] = {&scene.originalMesh(), &scene.innerShell(), &scene.outerShell(), &scene.additionalMesh()};
    std::vector<char> weights;
    const double buildMs = loadScene(r, p, scene, weights);
    const int64_t wbytes = static_cast<int64_t>(weights.size());
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
