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

#include <chrono>
#include <cstring>
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
