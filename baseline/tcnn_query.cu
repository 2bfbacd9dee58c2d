// baseline/tcnn_query.cu - runs the REFERENCE's neural query: tiny-cuda-nn's own
// NetworkWithInputEncoding<__half>::inference_mixed_precision (ext/tiny-cuda-nn include/tiny-cuda-nn/
// network_with_input_encoding.h:60-64, object.h:166-209), configured and loaded exactly as the renderer does
// (src/cuda_renderer_neural.cu:1647-1688 network JSON, :1752-1808 weight block [mlp | hg x n_points], :2024-2031 call).
// It is the arbiter for "neural outputs within fp16 tolerance" and the timed tcnn baseline of the query stage.
// Test/bench infrastructure only: built by baseline/Makefile into baseline/_ref/, never linked by the product.
//
//   tcnn_query <log2T> <base> <levels> <features> <scale> <points> <sh_degree|0> <width> <hidden>
//              <weights.bin> <swap_order 0|1> <inputs.f32> <n> <out.f16> [time_reps]
//
// weights.bin: the renderer's checkpoint file, raw fp16 [hg | mlp] (swap_order 1, the reference default) or [mlp | hg].
// inputs.f32 : n rows of input_dims fp32 (ray-major = column-major [input_dims x n], as networkInputs_).
// out.f16    : n rows of padded_output_width (16) fp16 (= outputs_).
// With time_reps > 0 it also prints one JSON line with the device time per call (CUDA events, default stream).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <string>
#include <vector>

#include <cuda_fp16.h>
#include <tiny-cuda-nn/common_device.h>
#include <tiny-cuda-nn/gpu_matrix.h>
#include <tiny-cuda-nn/network_with_input_encoding.h>

#define CK(x)                                                                                   \
    do {                                                                                        \
        cudaError_t e__ = (x);                                                                  \
        if (e__ != cudaSuccess) {                                                               \
            std::fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e__), __FILE__, __LINE__); \
            std::exit(2);                                                                       \
        }                                                                                       \
    } while (0)

static std::vector<char> slurp(const char* path) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) { std::fprintf(stderr, "cannot open %s\n", path); std::exit(2); }
    std::vector<char> b(static_cast<size_t>(f.tellg()));
    f.seekg(0);
    f.read(b.data(), static_cast<std::streamsize>(b.size()));
    return b;
}

int main(int argc, char** argv) {
    if (argc < 15) {
        std::fprintf(stderr, "usage: see header of baseline/tcnn_query.cu\n");
        return 2;
    }
    const int log2T = std::atoi(argv[1]), base = std::atoi(argv[2]), levels = std::atoi(argv[3]), feats = std::atoi(argv[4]);
    const float scale = static_cast<float>(std::atof(argv[5]));
    const uint32_t points = static_cast<uint32_t>(std::atoi(argv[6]));
    const int sh = std::atoi(argv[7]), width = std::atoi(argv[8]), hidden = std::atoi(argv[9]);
    const char* wpath = argv[10];
    const bool swapOrder = std::atoi(argv[11]) != 0;
    const char* ipath = argv[12];
    const size_t n = static_cast<size_t>(std::atoll(argv[13]));
    const char* opath = argv[14];
    const int reps = argc > 15 ? std::atoi(argv[15]) : 0;

    const uint32_t inputDims = points * 3u + (sh > 0 ? 3u : 0u);
    tcnn::json hashgridConfig = {
        {"otype", "HashGrid"}, {"n_dims_to_encode", 3}, {"n_levels", levels}, {"n_features_per_level", feats},
        {"log2_hashmap_size", log2T}, {"base_resolution", base}, {"per_level_scale", scale}, {"fixed_point_pos", false},
    };
    tcnn::json nested = tcnn::json::array();
    for (uint32_t p = 0; p < points; ++p) nested.push_back(hashgridConfig);
    if (sh > 0) nested.push_back({{"otype", "SphericalHarmonics"}, {"n_dims_to_encode", 3}, {"degree", sh}});
    tcnn::json encodingConfig = {{"otype", "Composite"}, {"nested", nested}};
    tcnn::json mlpConfig = {
        {"otype", "FullyFusedMLP"}, {"activation", "LeakyReLU"}, {"output_activation", "None"},
        {"n_neurons", width}, {"n_hidden_layers", hidden},
    };
    auto network = std::make_shared<tcnn::NetworkWithInputEncoding<__half>>(inputDims, 5u, encodingConfig, mlpConfig);
    const uint32_t outDims = network->padded_output_width();

    // loadWeightsFromFile, :1752-1808
    const size_t nParams = network->n_params();
    const size_t encN = network->encoding()->n_params(), mlpN = nParams - encN, hgN = encN / points;
    std::vector<char> file = slurp(wpath);
    if (file.size() != (mlpN + hgN) * sizeof(__half)) {
        std::fprintf(stderr, "weights size mismatch: got %zu, expected %zu (mlp %zu + hg %zu halfs)\n", file.size(),
                     (mlpN + hgN) * sizeof(__half), mlpN, hgN);
        return 3;
    }
    __half* params = nullptr;
    CK(cudaMalloc(&params, nParams * sizeof(__half)));
    const __half* src = reinterpret_cast<const __half*>(file.data());
    const __half* srcMlp = swapOrder ? src + hgN : src;
    const __half* srcHg = swapOrder ? src : src + mlpN;
    CK(cudaMemcpy(params, srcMlp, mlpN * sizeof(__half), cudaMemcpyHostToDevice));
    for (uint32_t i = 0; i < points; ++i)
        CK(cudaMemcpy(params + mlpN + i * hgN, srcHg, hgN * sizeof(__half), cudaMemcpyHostToDevice));
    network->set_params(params, params, nullptr);

    // networkInputs_ padded to the batch granularity with zero rows (:2017-2022)
    const size_t gran = tcnn::BATCH_SIZE_GRANULARITY;
    const size_t npad = (n + gran - 1) / gran * gran;
    std::vector<char> in;
    if (std::strcmp(ipath, "@rand") == 0) {  // timing runs: seeded uniform inputs made here instead of a large file
        in.resize(n * inputDims * sizeof(float));
        float* x = reinterpret_cast<float*>(in.data());
        uint32_t st = 99u;
        for (size_t i = 0; i < n * inputDims; ++i) {
            st = st * 1664525u + 1013904223u;
            x[i] = static_cast<float>(st >> 8) * (1.0f / 16777216.0f);
        }
    } else {
        in = slurp(ipath);
    }
    if (in.size() != n * inputDims * sizeof(float)) {
        std::fprintf(stderr, "inputs size mismatch: got %zu, expected %zu\n", in.size(), n * inputDims * sizeof(float));
        return 3;
    }
    float* dIn = nullptr;
    __half* dOut = nullptr;
    CK(cudaMalloc(&dIn, npad * inputDims * sizeof(float)));
    CK(cudaMemset(dIn, 0, npad * inputDims * sizeof(float)));
    CK(cudaMemcpy(dIn, in.data(), in.size(), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dOut, npad * outDims * sizeof(__half)));
    tcnn::GPUMatrix<float> inputMatrix(dIn, inputDims, static_cast<uint32_t>(npad));
    tcnn::GPUMatrix<__half> outputMatrix(dOut, outDims, static_cast<uint32_t>(npad));
    network->inference_mixed_precision(nullptr, inputMatrix, outputMatrix);
    CK(cudaDeviceSynchronize());
    std::vector<__half> out(n * outDims);
    CK(cudaMemcpy(out.data(), dOut, out.size() * sizeof(__half), cudaMemcpyDeviceToHost));
    if (std::strcmp(opath, "@none") != 0) {
        std::ofstream of(opath, std::ios::binary);
        of.write(reinterpret_cast<const char*>(out.data()), static_cast<std::streamsize>(out.size() * sizeof(__half)));
        of.close();
    }

    if (reps > 0) {
        cudaEvent_t a, b;
        CK(cudaEventCreate(&a));
        CK(cudaEventCreate(&b));
        for (int i = 0; i < 3; ++i) network->inference_mixed_precision(nullptr, inputMatrix, outputMatrix);
        CK(cudaEventRecord(a, 0));
        for (int i = 0; i < reps; ++i) network->inference_mixed_precision(nullptr, inputMatrix, outputMatrix);
        CK(cudaEventRecord(b, 0));
        CK(cudaEventSynchronize(b));
        float ms = 0.0f;
        CK(cudaEventElapsedTime(&ms, a, b));
        std::printf("{\"impl\": \"tcnn\", \"n\": %zu, \"input_dims\": %u, \"width\": %d, \"hidden\": %d, \"levels\": %d, "
                    "\"ms_per_call\": %.6f, \"gqueries_per_s\": %.4f, \"n_params\": %zu}\n",
                    npad, inputDims, width, hidden, levels, ms / reps, npad / (ms / reps) * 1e-6, nParams);
    }
    std::printf("tcnn_query: n=%zu (padded %zu) input_dims=%u out_dims=%u params=%zu\n", n, npad, inputDims, outDims, nParams);
    return 0;
}
