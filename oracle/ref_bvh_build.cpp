// oracle/ref_bvh_build.cpp — TEST INFRASTRUCTURE, built only into oracle/_ref/.
//
// Runs the reference's own Mesh::buildBvh (/root/reference/src/mesh_bvh.cpp:27-162, which is
// compiled from where it lies, together with the vendored ext/bvh builder) and hands back the
// flattened BvhNode[] and the leaf-ordered index buffer, so tests can feed the SAME built BVH to
// the reference headers, the restatement and the CUDA path.  Also times the
// ext/bvh CPU traversal (bvh::v2::Bvh::intersect with PrecomputedTri, the configuration of
// ext/bvh/test/benchmark.cpp) as the timed CPU baseline of config 1.
#define private public  // Mesh::bvhNodes_ is private; this is a harness TU, not product code
#include "mesh.h"
#undef private

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <tuple>
#include <thread>
#include <vector>

#include <bvh/v2/bvh.h>
#include <bvh/v2/default_builder.h>
#include <bvh/v2/executor.h>
#include <bvh/v2/ray.h>
#include <bvh/v2/stack.h>
#include <bvh/v2/thread_pool.h>
#include <bvh/v2/tri.h>
#include <bvh/v2/vec.h>

// Mesh's special members live in mesh.cu (CUDA); the harness never uploads anything.
Mesh::~Mesh() {}

extern "C" int ref_build_bvh(const float* vertices, int num_vertices, uint32_t* indices, int num_triangles,
                             void* nodes_out /* 44 B each, 2*num_triangles */, int* num_nodes,
                             float bounds_min[3], float bounds_max[3]) {
    Mesh m;
    m.vertices_.resize(num_vertices);
    std::memcpy(static_cast<void*>(m.vertices_.data()), vertices, sizeof(float) * 3 * num_vertices);
    m.indices_.resize(num_triangles);
    std::memcpy(m.indices_.data(), indices, sizeof(uint32_t) * 3 * num_triangles);
    m.buildBvh();
    std::memcpy(indices, m.indices_.data(), sizeof(uint32_t) * 3 * num_triangles);
    *num_nodes = static_cast<int>(m.bvhNodes_.size());
    std::memcpy(nodes_out, m.bvhNodes_.data(), sizeof(BvhNode) * m.bvhNodes_.size());
    Vec3 a = m.boundsMin(), b = m.boundsMax();
    bounds_min[0] = a.x; bounds_min[1] = a.y; bounds_min[2] = a.z;
    bounds_max[0] = b.x; bounds_max[1] = b.y; bounds_max[2] = b.z;
    return 0;
}

// ---- ext/bvh CPU traversal (BASELINE config 1; SURVEY.md 8c last row, 8d) ------------------------------------------
// bvh::v2::Bvh::intersect<false, false> + PrecomputedTri::intersect exactly as ext/bvh/test/benchmark.cpp and
// simple_example.cpp configure them (ext/bvh/src/bvh/v2/bvh.h:124-156, tri.h:58-74): DefaultBuilder Quality::High,
// primitives permuted into leaf order, non-robust traversal, SmallStack of 64, closest hit by shrinking ray.tmax.
// Rays are traced by `threads` std::threads over contiguous row blocks; the median wall time of `reps` passes is
// returned.  prim_out is the SOURCE triangle id (bvh.prim_ids), so it can be compared with any other tree's answer.
// This triangle test differs from the renderer's intersectTriangleIndexed (tolerance -eps on the barycentrics, no
// determinant cut, no facing cull), so ids agree except on edges and grazing hits: a cross-check and the timed CPU
// baseline, not the bit-exact arbiter.
namespace {
using Scalar = float;
using BVec3 = bvh::v2::Vec<Scalar, 3>;
using BBox = bvh::v2::BBox<Scalar, 3>;
using BTri = bvh::v2::Tri<Scalar, 3>;
using BNode = bvh::v2::Node<Scalar, 3>;
using BBvh = bvh::v2::Bvh<BNode>;
using BRay = bvh::v2::Ray<Scalar, 3>;
using PrecomputedTri = bvh::v2::PrecomputedTri<Scalar>;
}  // namespace

extern "C" int ref_extbvh_trace(const float* vertices, int num_vertices, const uint32_t* indices, int num_triangles,
                                const float* origins, const float* directions, size_t n, int threads, int reps, float tmin,
                                uint32_t* prim_out, float* t_out, float* uv_out /* 2n or null */,
                                float* normal_out /* 3n or null: unit geometric normal of the hit triangle */,
                                double* build_ms, double* trace_ms, int* threads_used) {
    (void)num_vertices;
    std::vector<BTri> tris(num_triangles);
    for (int i = 0; i < num_triangles; ++i) {
        const float* a = vertices + 3 * indices[3 * i], *b = vertices + 3 * indices[3 * i + 1], *c = vertices + 3 * indices[3 * i + 2];
        tris[i] = BTri(BVec3(a[0], a[1], a[2]), BVec3(b[0], b[1], b[2]), BVec3(c[0], c[1], c[2]));
    }
    const auto b0 = std::chrono::steady_clock::now();
    bvh::v2::ThreadPool pool;
    bvh::v2::ParallelExecutor executor(pool);
    std::vector<BBox> bboxes(tris.size());
    std::vector<BVec3> centers(tris.size());
    executor.for_each(0, tris.size(), [&](size_t begin, size_t end) {
        for (size_t i = begin; i < end; ++i) {
            bboxes[i] = tris[i].get_bbox();
            centers[i] = tris[i].get_center();
        }
    });
    typename bvh::v2::DefaultBuilder<BNode>::Config config;
    config.quality = bvh::v2::DefaultBuilder<BNode>::Quality::High;
    BBvh bvh = bvh::v2::DefaultBuilder<BNode>::build(pool, bboxes, centers, config);
    std::vector<PrecomputedTri> pre(tris.size());
    executor.for_each(0, tris.size(), [&](size_t begin, size_t end) {
        for (size_t i = begin; i < end; ++i) pre[i] = tris[bvh.prim_ids[i]];  // permuted primitives
    });
    if (build_ms) *build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - b0).count();

    int nt = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    if (nt < 1) nt = 1;
    if (threads_used) *threads_used = nt;
    constexpr size_t kInvalid = static_cast<size_t>(-1);
    auto work = [&](size_t begin, size_t end) {
        for (size_t i = begin; i < end; ++i) {
            BRay ray(BVec3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
                     BVec3(directions[3 * i], directions[3 * i + 1], directions[3 * i + 2]), tmin);
            size_t prim = kInvalid;
            Scalar u = 0, v = 0;
            bvh::v2::SmallStack<BBvh::Index, 64> stack;
            bvh.intersect<false, false>(ray, bvh.get_root().index, stack, [&](size_t lb, size_t le) {
                for (size_t k = lb; k < le; ++k)
                    if (auto hit = pre[k].intersect(ray)) {
                        prim = k;
                        std::tie(ray.tmax, u, v) = *hit;
                    }
                return prim != kInvalid;
            });
            const bool hit = prim != kInvalid;
            prim_out[i] = hit ? static_cast<uint32_t>(bvh.prim_ids[prim]) : 0xffffffffu;
            t_out[i] = hit ? ray.tmax : 1e30f;
            if (uv_out) { uv_out[2 * i] = u; uv_out[2 * i + 1] = v; }
            if (normal_out) {
                float nx = 0, ny = 0, nz = 0;
                if (hit) {
                    // PrecomputedTri's normal is cross(p0 - p1, p2 - p0) (left-handed): negate for the CCW normal
                    const auto& nn = pre[prim].n;
                    const float len = std::sqrt(nn[0] * nn[0] + nn[1] * nn[1] + nn[2] * nn[2]);
                    if (len > 0) { nx = -nn[0] / len; ny = -nn[1] / len; nz = -nn[2] / len; }
                }
                normal_out[3 * i] = nx; normal_out[3 * i + 1] = ny; normal_out[3 * i + 2] = nz;
            }
        }
    };
    std::vector<double> times;
    for (int r = 0; r < (reps > 0 ? reps : 1); ++r) {
        const auto t0 = std::chrono::steady_clock::now();
        std::vector<std::thread> pool2;
        const size_t chunk = (n + nt - 1) / nt;
        for (int k = 0; k < nt; ++k) {
            const size_t b = k * chunk, e = b + chunk < n ? b + chunk : n;
            if (b < e) pool2.emplace_back(work, b, e);
        }
        for (auto& th : pool2) th.join();
        times.push_back(std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    }
    std::sort(times.begin(), times.end());
    if (trace_ms) *trace_ms = times[times.size() / 2];
    return 0;
}
