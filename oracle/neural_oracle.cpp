// oracle/neural_oracle.cpp — TEST INFRASTRUCTURE (CPU oracle).  See neural_oracle.h.
#include "neural_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

// ---------------------------------------------------------------------------------------------
// fp16 emulation
// ---------------------------------------------------------------------------------------------
float orc_half_to_float(uint16_t h) {
    uint32_t sign = (h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1f;
    uint32_t man = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) {
            bits = sign;
        } else {
            int e = -1;
            do {
                man <<= 1;
                ++e;
            } while ((man & 0x400u) == 0);
            bits = sign | ((127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) {
        bits = sign | 0x7f800000u | (man << 13);
    } else {
        bits = sign | ((exp + 112) << 23) | (man << 13);
    }
    float f;
    std::memcpy(&f, &bits, 4);
    return f;
}

// Round (sign, mag * 2^-48) to fp16, round-to-nearest-even.  Exact for any product/sum of halfs.
static uint16_t round_scaled_to_half(bool neg, unsigned __int128 m) {
    uint16_t sign = neg ? 0x8000u : 0;
    if (m == 0) return sign;
    int k = 127;
    while (!((m >> k) & 1)) --k;
    int shift = k - 10 > 24 ? k - 10 : 24;
    unsigned __int128 q = m >> shift;
    unsigned __int128 rem = m & ((static_cast<unsigned __int128>(1) << shift) - 1);
    unsigned __int128 half = static_cast<unsigned __int128>(1) << (shift - 1);
    if (rem > half || (rem == half && (q & 1))) ++q;
    uint32_t qi = static_cast<uint32_t>(q);
    if (shift == 24) return static_cast<uint16_t>(sign | qi);  // subnormal, or exponent field 1
    int exp = shift - 23;
    if (qi == 2048) {
        qi = 1024;
        ++exp;
    }
    if (exp >= 31) return static_cast<uint16_t>(sign | 0x7c00u);
    return static_cast<uint16_t>(sign | (exp << 10) | (qi - 1024));
}

// Finite half -> signed integer in units of 2^-24.
static inline int64_t half_units(uint16_t h) {
    int exp = (h >> 10) & 0x1f;
    int64_t man = h & 0x3ff;
    int64_t v = exp == 0 ? man : ((man | 0x400) << (exp - 1));
    return (h & 0x8000u) ? -v : v;
}
static inline bool half_is_special(uint16_t h) { return ((h >> 10) & 0x1f) == 31; }

uint16_t orc_hfma(uint16_t a, uint16_t b, uint16_t c) {
    if (half_is_special(a) || half_is_special(b) || half_is_special(c)) {
        float r = std::fma(orc_half_to_float(a), orc_half_to_float(b), orc_half_to_float(c));
        return orc_float_to_half(r);
    }
    __int128 p = static_cast<__int128>(half_units(a)) * half_units(b);  // units 2^-48
    __int128 s = p + (static_cast<__int128>(half_units(c)) << 24);
    if (s == 0) {
        // IEEE sign of an exact zero sum: -0 only if both addends are -0 (RN mode).
        bool pneg = ((a ^ b) & 0x8000u) != 0;
        bool cneg = (c & 0x8000u) != 0;
        bool pzero = p == 0;
        bool czero = half_units(c) == 0;
        bool neg = (pzero && czero) ? (pneg && cneg) : false;
        return neg ? 0x8000u : 0;
    }
    bool neg = s < 0;
    return round_scaled_to_half(neg, static_cast<unsigned __int128>(neg ? -s : s));
}

uint16_t orc_hmul(uint16_t a, uint16_t b) {
    if (half_is_special(a) || half_is_special(b)) {
        return orc_float_to_half(orc_half_to_float(a) * orc_half_to_float(b));
    }
    __int128 p = static_cast<__int128>(half_units(a)) * half_units(b);
    bool neg = ((a ^ b) & 0x8000u) != 0;
    if (p == 0) return neg ? 0x8000u : 0;
    return round_scaled_to_half(neg, static_cast<unsigned __int128>(p < 0 ? -p : p));
}

uint16_t orc_float_to_half(float f) {
    uint32_t x;
    std::memcpy(&x, &f, 4);
    uint16_t sign = static_cast<uint16_t>((x >> 16) & 0x8000u);
    uint32_t ax = x & 0x7fffffffu;
    if (ax >= 0x7f800000u) return static_cast<uint16_t>(sign | 0x7c00u | (ax > 0x7f800000u ? 0x200u : 0));
    if (ax >= 0x477ff000u) return static_cast<uint16_t>(sign | 0x7c00u);  // >= 65520 rounds to inf
    if (ax < 0x33000001u) return sign;                                   // <= 2^-25 rounds to zero
    int e = static_cast<int>(ax >> 23) - 127;
    uint32_t man = (ax & 0x7fffffu) | 0x800000u;  // 24-bit significand, value = man * 2^(e-23)
    int shift = e >= -14 ? 13 : (13 + (-14 - e));  // bits to drop to reach fp16 quantum
    uint32_t q = man >> shift;
    uint32_t rem = man & ((1u << shift) - 1);
    uint32_t half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1))) ++q;
    if (e >= -14) {
        uint32_t bits = (static_cast<uint32_t>(e + 15) << 10) + (q - 1024);  // carry flows into exponent
        return static_cast<uint16_t>(sign | bits);
    }
    return static_cast<uint16_t>(sign | q);
}

// ---------------------------------------------------------------------------------------------
// Network configuration (RendererNeural ctor, cuda_renderer_neural.cu:1647-1688;
// GridEncodingTemplated ctor, encodings/grid.h:673-737)
// ---------------------------------------------------------------------------------------------
static uint32_t powi_u32(uint32_t b, int e) {
    uint32_t r = 1;
    for (int i = 0; i < e; ++i) r *= b;
    return r;
}

void OrcNetwork::configure(const nrb_nn_config* c) {
    nrb_nn_config d{};
    if (c) d = *c;
    if (d.log2_hashmap_size <= 0) d.log2_hashmap_size = 14;
    if (d.base_resolution <= 0) d.base_resolution = 16;
    if (d.n_points <= 0) d.n_points = 3;
    if (d.n_levels <= 0) d.n_levels = 8;
    if (d.n_features_per_level <= 0) d.n_features_per_level = 4;
    if (d.per_level_scale <= 0.0f) d.per_level_scale = 2.0f;
    if (d.sh_degree == 0) d.sh_degree = 4;       // 0 = reference value
    else if (d.sh_degree < 0) d.sh_degree = 0;   // negative = no direction encoding
    if (d.mlp_width <= 0) d.mlp_width = 128;
    if (d.n_hidden_layers <= 0) d.n_hidden_layers = 4;
    cfg = d;
    offsets.assign(cfg.n_levels + 1, 0);
    scales.assign(cfg.n_levels, 0.0f);
    resolutions.assign(cfg.n_levels, 0);
    const float log2_scale = std::log2(cfg.per_level_scale);
    uint32_t offset = 0;
    for (int l = 0; l < cfg.n_levels; ++l) {
        float scale = exp2f(l * log2_scale) * cfg.base_resolution - 1.0f;  // grid_scale
        uint32_t res = static_cast<uint32_t>(ceilf(scale)) + 1;            // grid_resolution
        scales[l] = scale;
        resolutions[l] = res;
        uint32_t max_params = std::numeric_limits<uint32_t>::max() / 2;
        uint32_t n = std::pow(static_cast<float>(res), 3) > static_cast<float>(max_params) ? max_params
                                                                                           : powi_u32(res, 3);
        n = (n + 7u) / 8u * 8u;                                      // next_multiple(.., 8)
        n = std::min(n, 1u << cfg.log2_hashmap_size);                // GridType::Hash
        offsets[l] = offset;
        offset += n;
    }
    offsets[cfg.n_levels] = offset;
    grid.assign(gridParams(), 0);
    mlp.assign(mlpParams(), 0);
}

// loadWeightsFromFile (cuda_renderer_neural.cu:1752-1808): [hg | mlp] when swapOrder, else [mlp | hg].
void OrcNetwork::load(const uint16_t* file, bool swapOrder) {
    const size_t ng = gridParams(), nm = mlpParams();
    const uint16_t* srcMlp = swapOrder ? file + ng : file;
    const uint16_t* srcHg = swapOrder ? file : file + nm;
    std::memcpy(mlp.data(), srcMlp, nm * 2);
    std::memcpy(grid.data(), srcHg, ng * 2);
    mlpF.resize(nm);
    for (size_t i = 0; i < nm; ++i) mlpF[i] = orc_half_to_float(mlp[i]);
}

// ---------------------------------------------------------------------------------------------
// Encoding: kernel_grid (encodings/grid.h:48-169) per point and level, then kernel_sh
// (encodings/spherical_harmonics.h:46-72, sh_enc common_device.h:476-502).  Row order follows
// CompositeEncoding::forward_impl (encodings/composite.h:215-258): point 0 levels.., point 1, ...
// then the SH block.
// ---------------------------------------------------------------------------------------------
static inline uint32_t grid_index_3d(uint32_t hashmap_size, uint32_t res, uint32_t x, uint32_t y, uint32_t z) {
    // grid_index (common_device.h:848-884) with coherent_prime_hash (:785-791)
    uint32_t stride = 1, index = 0;
    if (res <= 0x659u) {
        index += x * stride; stride *= res;
        index += y * stride; stride *= res;
        index += z * stride; stride *= res;
    } else {
        stride = 0xFFFFFFFFu;
    }
    if (hashmap_size < stride) index = (x * 1u) ^ (y * 2654435761u) ^ (z * 805459861u);
    return index % hashmap_size;
}

void OrcNetwork::encode(const float* in, uint16_t* enc) const {
    const int F = cfg.n_features_per_level;
    int row = 0;
    for (int p = 0; p < cfg.n_points; ++p) {
        const float* x = in + 3 * p;
        for (int l = 0; l < cfg.n_levels; ++l) {
            const uint16_t* table = grid.data() + static_cast<size_t>(offsets[l]) * F;
            const uint32_t hashmap_size = offsets[l + 1] - offsets[l];
            const float scale = scales[l];
            const uint32_t res = resolutions[l];
            float pos[3];
            uint32_t pg[3];
            for (int d = 0; d < 3; ++d) {  // pos_fract, common_device.h:1017-1029
                float v = fmaf(scale, x[d], 0.5f);
                float t = floorf(v);
                pg[d] = static_cast<uint32_t>(static_cast<int>(t));
                pos[d] = v - t;
            }
            uint16_t result[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (uint32_t idx = 0; idx < 8; ++idx) {
                float weight = 1;
                uint32_t c[3];
                for (int d = 0; d < 3; ++d) {
                    if ((idx & (1u << d)) == 0) {
                        weight *= 1 - pos[d];
                        c[d] = pg[d];
                    } else {
                        weight *= pos[d];
                        c[d] = pg[d] + 1;
                    }
                }
                const uint16_t w = orc_float_to_half(weight);
                const uint16_t* val = table + static_cast<size_t>(grid_index_3d(hashmap_size, res, c[0], c[1], c[2])) * F;
                for (int f = 0; f < F; ++f) result[f] = orc_hfma(w, val[f], result[f]);
            }
            for (int f = 0; f < F; ++f) enc[row++] = result[f];
        }
    }
    if (cfg.sh_degree > 0) {
        const float* d = in + 3 * cfg.n_points;
        float x = d[0] * 2.f - 1.f, y = d[1] * 2.f - 1.f, z = d[2] * 2.f - 1.f;
        float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
        float o[16];
        int deg = cfg.sh_degree;
        o[0] = 0.28209479177387814f;
        if (deg > 1) {
            o[1] = -0.48860251190291987f * y;
            o[2] = 0.48860251190291987f * z;
            o[3] = -0.48860251190291987f * x;
        }
        if (deg > 2) {
            o[4] = 1.0925484305920792f * xy;
            o[5] = -1.0925484305920792f * yz;
            o[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
            o[7] = -1.0925484305920792f * xz;
            o[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
        }
        if (deg > 3) {
            o[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
            o[10] = 2.8906114426405538f * xy * z;
            o[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
            o[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
            o[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
            o[14] = 1.4453057213202769f * z * (x2 - y2);
            o[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
        }
        for (int i = 0; i < deg * deg && i < 16; ++i) enc[row++] = orc_float_to_half(o[i]);
    }
}

// kernel_mlp_fused (src/fully_fused_mlp.cu:499-557): weights row-major [out][in] (:659-671),
// LeakyReLU slope 0.01 applied in fp16 (common_device.h:123-129), no output activation.
void OrcNetwork::forward(const float* inputs, int n, uint16_t* out) const {
    const int E = encodedDims(), W = cfg.mlp_width, D = inputDims();
    const uint16_t slope = orc_float_to_half(0.01f);
    std::vector<uint16_t> enc(E);
    std::vector<float> a(std::max(E, W)), b(W);
    const std::vector<float>& wf = mlpF;
    for (int q = 0; q < n; ++q) {
        encode(inputs + static_cast<size_t>(q) * D, enc.data());
        for (int i = 0; i < E; ++i) a[i] = orc_half_to_float(enc[i]);
        const float* w = wf.data();
        int in_w = E;
        for (int layer = 0; layer < cfg.n_hidden_layers; ++layer) {
            for (int o = 0; o < W; ++o) {
                double acc = 0.0;
                const float* wr = w + static_cast<size_t>(o) * in_w;
                for (int k = 0; k < in_w; ++k) acc += static_cast<double>(wr[k]) * a[k];
                uint16_t h = orc_float_to_half(static_cast<float>(acc));
                bool pos = !(h & 0x8000u) && (h & 0x7fffu) != 0;
                if (!pos) h = orc_hmul(h, slope);
                b[o] = orc_half_to_float(h);
            }
            w += static_cast<size_t>(W) * in_w;
            in_w = W;
            std::swap(a, b);
            if (static_cast<int>(b.size()) < W) b.resize(W);
        }
        for (int o = 0; o < 16; ++o) {
            double acc = 0.0;
            const float* wr = w + static_cast<size_t>(o) * W;
            for (int k = 0; k < W; ++k) acc += static_cast<double>(wr[k]) * a[k];
            out[static_cast<size_t>(q) * 16 + o] = orc_float_to_half(static_cast<float>(acc));
        }
    }
}

// C entry points of the software fp16 helpers (tests/test_oracle_pinning.py checks them against numpy).
extern "C" {
uint16_t orc_c_float_to_half(float f) { return orc_float_to_half(f); }
float orc_c_half_to_float(uint16_t h) { return orc_half_to_float(h); }
uint16_t orc_c_hfma(uint16_t a, uint16_t b, uint16_t c) { return orc_hfma(a, b, c); }
}
