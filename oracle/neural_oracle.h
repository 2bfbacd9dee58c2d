// oracle/neural_oracle.h — TEST INFRASTRUCTURE (CPU oracle), never part of the product path.
//
// CPU restatement of the arithmetic of the reference's neural query, which lives in the vendored
// tiny-cuda-nn (julcst fork @2a4340e under /root/reference/ext/tiny-cuda-nn) and is CUDA-only:
//   NetworkWithInputEncoding<__half> = Composite(n_points x HashGrid + SphericalHarmonics)
//   -> FullyFusedMLP (LeakyReLU 0.01, no output activation, no biases).
// fp16 is emulated bit-exactly for the encodings (the grid's __hfma chain, the fp32->fp16 casts);
// the MLP accumulates in double and rounds to fp16 once per neuron, because the reference's HMMA
// fp16-accumulate order is not reproducible.  PARITY UNPINNED: no runnable
// reference for this function exists in this container; external pins are the offset-table
// known answers of ext/tiny-cuda-nn/tests/test_grid.cu:37-92 (tests/test_oracle_neural.py).
#pragma once

#include <cstddef>
#include <cstdint>
#include <vector>

#include "../include/nrb200.h"

float orc_half_to_float(uint16_t h);
uint16_t orc_float_to_half(float f);                      // round-to-nearest-even (__float2half_rn)
uint16_t orc_hfma(uint16_t a, uint16_t b, uint16_t c);    // a*b+c, single rounding (__hfma)
uint16_t orc_hmul(uint16_t a, uint16_t b);

struct OrcNetwork {
    nrb_nn_config cfg{};
    std::vector<uint32_t> offsets;   // per-level entry offsets, size n_levels+1 (grid.h:699-730)
    std::vector<float> scales;       // grid_scale per level (common_device.h:886-891)
    std::vector<uint32_t> resolutions;
    std::vector<uint16_t> grid;      // one shared table (the reference tiles it n_points times)
    std::vector<float> mlpF;         // mlp as fp32 (exact)
    std::vector<uint16_t> mlp;       // W0 [width][enc], W1.. [width][width], Wlast [16][width]

    void configure(const nrb_nn_config* c);
    int shCoeffs() const { return cfg.sh_degree * cfg.sh_degree; }
    int inputDims() const { return cfg.n_points * 3 + (cfg.sh_degree > 0 ? 3 : 0); }
    int encodedDims() const { return cfg.n_points * cfg.n_levels * cfg.n_features_per_level + shCoeffs(); }
    size_t gridParams() const { return static_cast<size_t>(offsets.back()) * cfg.n_features_per_level; }
    size_t mlpParams() const {
        size_t w = cfg.mlp_width;
        return w * encodedDims() + static_cast<size_t>(cfg.n_hidden_layers - 1) * w * w + 16 * w;
    }
    size_t fileBytes() const { return (gridParams() + mlpParams()) * 2; }
    size_t paramsBytes() const { return (gridParams() * cfg.n_points + mlpParams()) * 2; }
    void load(const uint16_t* file, bool swapOrder);
    // Encoded network input of one query (fp16 bits), encodedDims() values, rows in Composite order.
    void encode(const float* in, uint16_t* enc) const;
    // inputs [n][inputDims] fp32 -> outputs [n][16] fp16 bits.
    void forward(const float* inputs, int n, uint16_t* out) const;
};
