// nrb_neural.cuh — device description of the neural query network.
//
// Replaces tiny-cuda-nn's NetworkWithInputEncoding<__half> on this path (reference
// src/cuda_renderer_neural.cu:1647-1702): Composite(n_points x HashGrid + SphericalHarmonics) ->
// fully fused MLP (LeakyReLU 0.01, no biases, 16 padded outputs).  The reference tiles ONE grid
// table into n_points identical copies (:1798-1804); here the table is stored once and shared.
#pragma once

#include <cuda_fp16.h>
#include <stdint.h>

namespace nrb {

constexpr int kMaxLevels = 16;

struct NetDev {
    const __half* grid;   // [offsets[L]][F]
    const __half* mlp;    // W0 [W][E] | W1..W(H-1) [W][W] | Wlast [16][W], row-major [out][in]
    int nPoints, nLevels, nFeatures, shDegree, width, nHidden;
    int inputDims;        // 3*nPoints (+3 when shDegree > 0)
    int encDims;          // nPoints*nLevels*nFeatures + shDegree^2
    int kernelChoice;     // NRB_OPT_QUERY_KERNEL of the owning renderer: low byte 0 auto, 1 mma.sync, 2 fused tcgen05, 3 split;
                          // higher bits = the experiment flags below
    int debug;            // experiments only: bit 0 = skip the table loads, bit 1 = skip the MLP,
                          // bit 2 = split encoder gathers from global memory, bit 3 = 1024-row chunks
    // split query (nrb_query_split.cu): weights pre-packed as UMMA operands, encoded-row scratch
    const void* packedW;  // split_packed_bytes(net) bytes, written by split_prepack()
    void* xt;             // [xtRows/128][EK/8][128][8] fp16 encoded tiles
    uint32_t xtRows;      // capacity of xt in rows (multiple of 128); 0 = split path unavailable
    int* encCursor;       // [nLevels] k_query_encode's per-level tile cursors (zero between launches)
    uint32_t levelOffset[kMaxLevels];
    uint32_t levelSize[kMaxLevels];     // hashmap_size of the level (entries)
    uint32_t levelRes[kMaxLevels];
    float levelScale[kMaxLevels];
    uint8_t levelDense[kMaxLevels];     // res^3 <= size: linear index, else coherent-prime hash
    uint8_t levelPow2[kMaxLevels];      // size is a power of two (mask instead of %)
};

// Enqueue the query for n rows: inputs fp32 [n][inputDims] (or, with dCount != nullptr, the row
// count is read on the device from *dCount and n is only the launch bound), outputs fp16 [n][16].
int launch_neural_query(const NetDev& net, const float* in, size_t n, const int* dCount, __half* out,
                        cudaStream_t s);
// The tcgen05/TMEM kernel (nrb_neural_tc.cu) and the kernel choice: 0 = automatic (split when the
// shape is supported and scratch is attached, else fused tcgen05, else mma.sync), 1 = mma.sync kernel,
// 2 = fused tcgen05 kernel, 3 = split encode + tcgen05 MLP.
bool neural_query_tc_supported(const NetDev& net);
int launch_neural_query_tc(const NetDev& net, const float* in, size_t n, const int* dCount, __half* out,
                           cudaStream_t s);
// The split query (nrb_query_split.cu): k_query_encode (one hash-grid level per CTA, the level's table
// resident in shared memory) -> encoded tiles in HBM/L2 -> k_query_mlp (tcgen05, two tiles in flight per
// SM, activations fed by bulk async copies).  Kernel choice 3 (and the automatic default when supported).
bool split_query_supported(const NetDev& net);
size_t split_packed_bytes(const NetDev& net);
size_t split_xt_bytes(const NetDev& net, size_t rows);           // rows rounded up to 128 inside
int split_prepack(const NetDev& net, void* packedW, cudaStream_t s);
int launch_split_query(const NetDev& net, const float* in, size_t n, const int* dCount, __half* out, cudaStream_t s);
// The same two kernels with the encoded tiles and cursors in a file-scope __device__ array of nrb_query_split.cu
// (nrb_neural_query_static): no allocation in the call, at most split_static_max_rows() rows.  < 0 = refused
// (-1 shape or size, -2 CUDA), *why says it.
int launch_split_query_static(const NetDev& net, const float* in, size_t n, __half* out, cudaStream_t s, const char** why);
size_t split_static_max_rows();
// Bytes of dynamic shared memory the query kernel needs for this network (0 = unsupported shape).
size_t neural_query_smem(const NetDev& net, int* warpsOut);

}  // namespace nrb
