// nrb_bvh_gpu.cu — BVH construction on the GPU: the step BEFORE the hot path.
//
// The reference builds its BVH on the CPU with ext/bvh's High-quality builder (Mesh::buildBvh,
// src/mesh_bvh.cpp:27-162: seconds for multi-million-triangle meshes) and so does nrb_build_bvh
// (nrb_bvh_build.cpp, binned SAH).  nrb_build_bvh_gpu produces the SAME output contract - BvhNode[]
// with child indices / leaf ranges, a leaf-ordered index buffer and the source triangle of every
// output triangle - in milliseconds, for dynamic scenes and large meshes:
//   1. per-triangle boxes + scene box (warp-reduced atomics on order-preserving integer keys)
//   2. 30-bit Morton code of the box centre, made unique as a 64-bit key (code << 32 | triangle)
//   3. radix sort of the keys (CUB device-wide sort: a library primitive, like the reference's use of
//      std::sort inside ext/bvh)
//   4. the binary radix tree over the sorted keys, one thread per inner node (Karras 2012)
//   5. bottom-up box fitting, the second thread to reach a node continues upward
//   6. subtrees of <= max_leaf triangles are folded into one leaf (their keys are contiguous)
// The tree is an LBVH, not the reference's SAH tree: primitive ids (= position in leaf order) differ
// from a reference build, so bit-exact id parity is defined on prim_ids (source triangles); t is
// produced by the same triangle test on the same triangles and is bit-identical.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cub/device/device_radix_sort.cuh>
#include <string>

#include "../../include/nrb200.h"

namespace {

struct Box {
    float lo[3], hi[3];
};

__device__ __forceinline__ int ordered(float f) {  // monotone float -> int
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float unordered(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

__global__ void k_lbvh_boxes(const float* __restrict__ v, const uint32_t* __restrict__ idx, int n, Box* boxes, int* scene) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
    if (i < n) {
        for (int c = 0; c < 3; ++c) {
            const float* p = v + 3 * static_cast<size_t>(idx[3 * static_cast<size_t>(i) + c]);
            for (int k = 0; k < 3; ++k) {
                lo[k] = fminf(lo[k], p[k]);
                hi[k] = fmaxf(hi[k], p[k]);
            }
        }
        Box b;
        for (int k = 0; k < 3; ++k) { b.lo[k] = lo[k]; b.hi[k] = hi[k]; }
        boxes[i] = b;
    }
    for (int k = 0; k < 3; ++k) {
        float a = lo[k], z = hi[k];
        for (int o = 16; o > 0; o >>= 1) {
            a = fminf(a, __shfl_xor_sync(0xffffffffu, a, o));
            z = fmaxf(z, __shfl_xor_sync(0xffffffffu, z, o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(scene + k, ordered(a));
            atomicMax(scene + 3 + k, ordered(z));
        }
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x) {  // 10 bits -> every third bit
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__global__ void k_lbvh_keys(const Box* __restrict__ boxes, int n, const int* __restrict__ scene, uint64_t* keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t q[3];
    for (int k = 0; k < 3; ++k) {
        const float lo = unordered(scene[k]), hi = unordered(scene[3 + k]);
        const float ext = hi - lo;
        const float c = 0.5f * (boxes[i].lo[k] + boxes[i].hi[k]);
        const float t = ext > 0.0f ? (c - lo) / ext : 0.0f;
        q[k] = static_cast<uint32_t>(fminf(fmaxf(t * 1024.0f, 0.0f), 1023.0f));
    }
    const uint32_t code = (spread10(q[0]) << 2) | (spread10(q[1]) << 1) | spread10(q[2]);
    keys[i] = (static_cast<uint64_t>(code) << 32) | static_cast<uint32_t>(i);
}

// Inner node i of the radix tree over n sorted unique keys: its key range and split (Karras 2012, §4).
__device__ __forceinline__ int common_prefix(const uint64_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    return __clzll(static_cast<long long>(keys[i] ^ keys[j]));
}

struct Inner {
    int left, right;   // >= 0 inner node, < 0 leaf ~index
    int first, last;   // key range
    int parent;
};

__global__ void k_lbvh_tree(const uint64_t* __restrict__ keys, int n, Inner* inner, int* leafParent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = common_prefix(keys, n, i, i + 1) > common_prefix(keys, n, i, i - 1) ? 1 : -1;
    const int dmin = common_prefix(keys, n, i, i - d);
    int lmax = 2;
    while (common_prefix(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (common_prefix(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = common_prefix(keys, n, i, j);
    int s = 0;
    for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
        if (common_prefix(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t == 1) break;
    }
    const int gamma = i + s * d + min(d, 0);
    const int first = min(i, j), last = max(i, j);
    const int left = first == gamma ? ~gamma : gamma;
    const int right = last == gamma + 1 ? ~(gamma + 1) : gamma + 1;
    inner[i].left = left;   // every field has exactly one writer: a node's own thread, or its parent's
    inner[i].right = right;
    inner[i].first = first;
    inner[i].last = last;
    if (left >= 0) inner[left].parent = i; else leafParent[~left] = i;
    if (right >= 0) inner[right].parent = i; else leafParent[~right] = i;
    if (i == 0) inner[0].parent = -1;
}

__device__ __forceinline__ void merge(Box& a, const Box& b) {
    for (int k = 0; k < 3; ++k) {
        a.lo[k] = fminf(a.lo[k], b.lo[k]);
        a.hi[k] = fmaxf(a.hi[k], b.hi[k]);
    }
}

// Output node numbering: inner node i -> i (root = 0), leaf (sorted position) j -> n - 1 + j.
__global__ void k_lbvh_fit(const uint64_t* __restrict__ keys, const Box* __restrict__ boxes, int n, const Inner* __restrict__ inner,
                           const int* __restrict__ leafParent, int* visits, Box* innerBox, nrb_bvh_node* out, int maxLeaf) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    Box b = boxes[static_cast<uint32_t>(keys[j])];
    nrb_bvh_node leaf{};
    for (int k = 0; k < 3; ++k) { leaf.bounds_min[k] = b.lo[k]; leaf.bounds_max[k] = b.hi[k]; }
    leaf.first = j;
    leaf.count = 1;
    leaf.is_leaf = 1;
    out[n - 1 + j] = leaf;
    int node = n > 1 ? leafParent[j] : -1;
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(visits + node, 1) == 0) return;  // the sibling subtree is not finished yet
        __threadfence();
        const Inner nd = inner[node];
        // child boxes: a leaf's is an input; an inner child's was written by another thread -> read past L1
        auto child_box = [&](int c) {
            Box r;
            if (c < 0) return boxes[static_cast<uint32_t>(keys[~c])];
            const float* p = reinterpret_cast<const float*>(innerBox + c);
            for (int k = 0; k < 3; ++k) { r.lo[k] = __ldcg(p + k); r.hi[k] = __ldcg(p + 3 + k); }
            return r;
        };
        Box u = child_box(nd.left);
        const Box w = child_box(nd.right);
        merge(u, w);
        innerBox[node] = u;
        nrb_bvh_node o{};
        for (int k = 0; k < 3; ++k) { o.bounds_min[k] = u.lo[k]; o.bounds_max[k] = u.hi[k]; }
        const int size = nd.last - nd.first + 1;
        if (size <= maxLeaf) {  // fold the subtree into one leaf (reachable only if the parent is larger)
            o.first = nd.first;
            o.count = size;
            o.is_leaf = 1;
        } else {
            o.left = nd.left >= 0 ? nd.left : n - 1 + ~nd.left;
            o.right = nd.right >= 0 ? nd.right : n - 1 + ~nd.right;
        }
        out[node] = o;
        node = nd.parent;
    }
}

__global__ void k_lbvh_permute(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ idx, int n, uint32_t* outIdx, uint32_t* prim) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint32_t src = static_cast<uint32_t>(keys[j]);
    prim[j] = src;
    for (int c = 0; c < 3; ++c) outIdx[3 * static_cast<size_t>(j) + c] = idx[3 * static_cast<size_t>(src) + c];
}

struct Scratch {
    void* p[12] = {};
    ~Scratch() { for (void* q : p) if (q) cudaFree(q); }
};

thread_local std::string g_err;

}  // namespace

extern "C" {

const char* nrb_build_bvh_gpu_error(void) { return g_err.c_str(); }

int nrb_build_bvh_gpu(const float* vertices, int num_vertices, uint32_t* indices, int num_triangles, nrb_bvh_node* nodes,
                      int* num_nodes, uint32_t* prim_ids, int device, int max_leaf, float* build_ms) {
    auto fail = [&](int code, const std::string& m) { g_err = m; return code; };
    if (!vertices || !indices || !nodes || !num_nodes || num_vertices <= 0 || num_triangles <= 0)
        return fail(NRB_ERR_INVALID, "null or empty mesh");
    if (max_leaf <= 0) max_leaf = 4;
    if (max_leaf > 32) max_leaf = 32;
    for (size_t i = 0; i < 3 * static_cast<size_t>(num_triangles); ++i)
        if (indices[i] >= static_cast<uint32_t>(num_vertices)) return fail(NRB_ERR_INVALID, "vertex index out of range");
#define LB_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return fail(NRB_ERR_CUDA, std::string(#x) + ": " + cudaGetErrorString(e_)); } while (0)
    if (device < 0) {
        if (cudaGetDevice(&device) != cudaSuccess) return fail(NRB_ERR_CUDA, "no usable CUDA device (the GPU builder has no CPU path)");
    }
    LB_CUDA(cudaSetDevice(device));
    const int n = num_triangles;
    Scratch s;
    float* dV; uint32_t *dIdx, *dOutIdx, *dPrim; Box *dBoxes, *dInnerBox; int *dScene, *dLeafParent, *dVisits;
    uint64_t *dKeys, *dKeysSorted; Inner* dInner; nrb_bvh_node* dOut;
    auto take = [&](int slot, size_t bytes, void** out) { cudaError_t e = cudaMalloc(out, bytes ? bytes : 16); s.p[slot] = *out; return e; };
    LB_CUDA(take(0, sizeof(float) * 3 * num_vertices, reinterpret_cast<void**>(&dV)));
    LB_CUDA(take(1, sizeof(uint32_t) * 3 * n, reinterpret_cast<void**>(&dIdx)));
    LB_CUDA(take(2, sizeof(uint32_t) * 3 * n, reinterpret_cast<void**>(&dOutIdx)));
    LB_CUDA(take(3, sizeof(uint32_t) * n, reinterpret_cast<void**>(&dPrim)));
    LB_CUDA(take(4, sizeof(Box) * n, reinterpret_cast<void**>(&dBoxes)));
    LB_CUDA(take(5, sizeof(Box) * n, reinterpret_cast<void**>(&dInnerBox)));
    LB_CUDA(take(6, sizeof(int) * (6 + 2 * static_cast<size_t>(n)), reinterpret_cast<void**>(&dScene)));
    dLeafParent = dScene + 6;
    dVisits = dLeafParent + n;
    LB_CUDA(take(7, sizeof(uint64_t) * n, reinterpret_cast<void**>(&dKeys)));
    LB_CUDA(take(8, sizeof(uint64_t) * n, reinterpret_cast<void**>(&dKeysSorted)));
    LB_CUDA(take(9, sizeof(Inner) * n, reinterpret_cast<void**>(&dInner)));
    LB_CUDA(take(10, sizeof(nrb_bvh_node) * (2 * static_cast<size_t>(n)), reinterpret_cast<void**>(&dOut)));
    size_t sortBytes = 0;
    LB_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, sortBytes, dKeys, dKeysSorted, n, 0, 62));
    void* dSort;
    LB_CUDA(take(11, sortBytes, &dSort));

    LB_CUDA(cudaMemcpy(dV, vertices, sizeof(float) * 3 * num_vertices, cudaMemcpyHostToDevice));
    LB_CUDA(cudaMemcpy(dIdx, indices, sizeof(uint32_t) * 3 * n, cudaMemcpyHostToDevice));
    cudaEvent_t e0, e1;
    LB_CUDA(cudaEventCreate(&e0));
    LB_CUDA(cudaEventCreate(&e1));
    const int h[6] = {0x7fffffff, 0x7fffffff, 0x7fffffff, static_cast<int>(0x80000000u), static_cast<int>(0x80000000u), static_cast<int>(0x80000000u)};
    LB_CUDA(cudaMemcpy(dScene, h, sizeof(h), cudaMemcpyHostToDevice));
    LB_CUDA(cudaEventRecord(e0));
    LB_CUDA(cudaMemsetAsync(dVisits, 0, sizeof(int) * n));
    const int T = 256, G = (n + T - 1) / T;
    k_lbvh_boxes<<<G, T>>>(dV, dIdx, n, dBoxes, dScene);
    k_lbvh_keys<<<G, T>>>(dBoxes, n, dScene, dKeys);
    LB_CUDA(cub::DeviceRadixSort::SortKeys(dSort, sortBytes, dKeys, dKeysSorted, n, 0, 62));
    if (n > 1) k_lbvh_tree<<<(n - 1 + T - 1) / T, T>>>(dKeysSorted, n, dInner, dLeafParent);
    k_lbvh_fit<<<G, T>>>(dKeysSorted, dBoxes, n, dInner, dLeafParent, dVisits, dInnerBox, dOut, max_leaf);
    k_lbvh_permute<<<G, T>>>(dKeysSorted, dIdx, n, dOutIdx, dPrim);
    LB_CUDA(cudaEventRecord(e1));
    LB_CUDA(cudaEventSynchronize(e1));
    LB_CUDA(cudaGetLastError());
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (build_ms) *build_ms = ms;
    const int total = n == 1 ? 1 : 2 * n - 1;
    if (n == 1) {  // the single leaf is the root
        LB_CUDA(cudaMemcpy(nodes, dOut + 0, sizeof(nrb_bvh_node), cudaMemcpyDeviceToHost));
    } else {
        LB_CUDA(cudaMemcpy(nodes, dOut, sizeof(nrb_bvh_node) * total, cudaMemcpyDeviceToHost));
    }
    LB_CUDA(cudaMemcpy(indices, dOutIdx, sizeof(uint32_t) * 3 * n, cudaMemcpyDeviceToHost));
    if (prim_ids) LB_CUDA(cudaMemcpy(prim_ids, dPrim, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost));
    *num_nodes = total;
#undef LB_CUDA
    return NRB_OK;
}

}  // extern "C"
