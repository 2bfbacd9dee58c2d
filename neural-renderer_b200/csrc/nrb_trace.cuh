// nrb_trace.cuh — closest-hit BVH traversal and hit-record construction.
//
// Same results as the reference's traceMeshWithMode (src/mesh_traversal.cuh:118-222) with a
// different machine mapping: both child boxes come from one 64-byte NodePair (4 x LDG.128), the
// near child is followed directly instead of being pushed and popped, the far child is pushed
// together with its entry distance so the reference's "re-test the popped box against the
// current best t" becomes a register compare, and candidate triangles are 3 x LDG.128
// TriBlocks with the facing normal pre-computed.  The visit ORDER of triangles is the
// reference's (near child first by the same rightNear < leftNear rule, leaf triangles
// ascending), so ties in t resolve to the same primitive id.
#pragma once

#include "nrb_scene.cuh"

namespace nrb {

constexpr int kTraceStack = 64;       // checked against the tree depth at upload
constexpr int kLeafCountBits = 5;     // leaf ref = ~((first << 5) | (count - 1))

__host__ __device__ __forceinline__ int leaf_ref(int first, int count) {
    return ~((first << kLeafCountBits) | (count - 1));
}

// intersectAabb, reference src/mesh_traversal.cuh:61-93 (same operations, same order).
__device__ __forceinline__ bool slab(V3 o, V3 inv, float mnx, float mny, float mnz, float mxx, float mxy,
                                     float mxz, float tMax, float& tNear) {
    // the reference's +-1e-10 padding of the box is applied once at upload (nrb_api.cu:relayout_mesh,
    // same fp32 subtraction/addition), so the boxes arriving here are already padded
    float t1 = (mnx - o.x) * inv.x, t2 = (mxx - o.x) * inv.x;
    float tmin = fminf(t1, t2), tmax = fmaxf(t1, t2);
    t1 = (mny - o.y) * inv.y; t2 = (mxy - o.y) * inv.y;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    t1 = (mnz - o.z) * inv.z; t2 = (mxz - o.z) * inv.z;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    tNear = tmin;
    return !(tmax < 0.0f || tmin > tMax || tmin > tmax);
}

// Both boxes of a NodePair against one ray with packed fp32x2 arithmetic (FADD2 / FMUL2 on sm_100: two
// IEEE round-to-nearest results per instruction, the same values as the scalar form above; min/max have no
// packed form).  The ray's origin and reciprocal direction are kept as (x, y) and (z, z) register couples.
// Returns the best-independent part of intersectAabb per box (tmax >= 0 and tmin <= tmax) and the entry
// distances; the caller adds "entry distance not beyond the current best".
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 submul2(f32x2 a, f32x2 o, f32x2 inv) {  // (a - o) * inv, two roundings
    f32x2 d, r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(o));
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(d), "l"(inv));
    return r;
}
struct RayPack {
    f32x2 oxy, ozz, ixy, izz;
};
__device__ __forceinline__ RayPack ray_pack(V3 o, V3 inv) {
    return RayPack{pack2(o.x, o.y), pack2(o.z, o.z), pack2(inv.x, inv.y), pack2(inv.z, inv.z)};
}
__device__ __forceinline__ bool slab_finish(float t1x, float t1y, float t1z, float t2x, float t2y, float t2z, float& tNear) {
    float tmin = fminf(t1x, t2x), tmax = fmaxf(t1x, t2x);
    tmin = fmaxf(tmin, fminf(t1y, t2y)); tmax = fminf(tmax, fmaxf(t1y, t2y));
    tmin = fmaxf(tmin, fminf(t1z, t2z)); tmax = fminf(tmax, fmaxf(t1z, t2z));
    tNear = tmin;
    return !(tmax < 0.0f || tmin > tmax);
}
__device__ __forceinline__ bool slab_finish_far(float t1x, float t1y, float t1z, float t2x, float t2y, float t2z, float& tNear,
                                                float& tFar) {
    float tmin = fminf(t1x, t2x), tmax = fmaxf(t1x, t2x);
    tmin = fmaxf(tmin, fminf(t1y, t2y)); tmax = fminf(tmax, fmaxf(t1y, t2y));
    tmin = fmaxf(tmin, fminf(t1z, t2z)); tmax = fminf(tmax, fmaxf(t1z, t2z));
    tNear = tmin;
    tFar = tmax;
    return !(tmax < 0.0f || tmin > tmax);
}
// slab_pair that also hands out the exit distances (the same arithmetic; trace_sync's lower bound needs them)
__device__ __forceinline__ void slab_pair_far(const RayPack& rp, float4 a, float4 b, float4 c, bool& okL, float& nearL, float& farL,
                                              bool& okR, float& nearR, float& farR) {
    float l1x, l1y, l2x, l2y, l1z, l2z, r1x, r1y, r2x, r2y, r1z, r2z;
    unpack2(submul2(pack2(a.x, a.y), rp.oxy, rp.ixy), l1x, l1y);
    unpack2(submul2(pack2(a.z, a.w), rp.oxy, rp.ixy), l2x, l2y);
    unpack2(submul2(pack2(b.x, b.y), rp.oxy, rp.ixy), r1x, r1y);
    unpack2(submul2(pack2(b.z, b.w), rp.oxy, rp.ixy), r2x, r2y);
    unpack2(submul2(pack2(c.x, c.y), rp.ozz, rp.izz), l1z, l2z);
    unpack2(submul2(pack2(c.z, c.w), rp.ozz, rp.izz), r1z, r2z);
    okL = slab_finish_far(l1x, l1y, l1z, l2x, l2y, l2z, nearL, farL);
    okR = slab_finish_far(r1x, r1y, r1z, r2x, r2y, r2z, nearR, farR);
}
__device__ __forceinline__ void slab_pair(const RayPack& rp, float4 a, float4 b, float4 c, bool& okL, float& nearL,
                                          bool& okR, float& nearR) {
    float l1x, l1y, l2x, l2y, l1z, l2z, r1x, r1y, r2x, r2y, r1z, r2z;
    unpack2(submul2(pack2(a.x, a.y), rp.oxy, rp.ixy), l1x, l1y);
    unpack2(submul2(pack2(a.z, a.w), rp.oxy, rp.ixy), l2x, l2y);
    unpack2(submul2(pack2(b.x, b.y), rp.oxy, rp.ixy), r1x, r1y);
    unpack2(submul2(pack2(b.z, b.w), rp.oxy, rp.ixy), r2x, r2y);
    unpack2(submul2(pack2(c.x, c.y), rp.ozz, rp.izz), l1z, l2z);
    unpack2(submul2(pack2(c.z, c.w), rp.ozz, rp.izz), r1z, r2z);
    okL = slab_finish(l1x, l1y, l1z, l2x, l2y, l2z, nearL);
    okR = slab_finish(r1x, r1y, r1z, r2x, r2y, r2z, nearR);
}

// Counters for the roofline's algorithmic bytes (node pairs fetched, triangles fetched).
struct TraceStats {
    unsigned int pairs, tris;
};

template <int MODE, bool STATS = false>
__device__ __forceinline__ Hit trace_closest(const MeshDev& m, V3 o, V3 d, TraceStats* st = nullptr) {
    Hit best{1e30f, 0.0f, 0.0f, 0u};
    if (m.numTriangles <= 0) return best;
    const V3 inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    float tn;
    if (!slab(o, inv, m.rootMin[0], m.rootMin[1], m.rootMin[2], m.rootMax[0], m.rootMax[1], m.rootMax[2], best.t, tn))
        return best;
    const RayPack rp = ray_pack(o, inv);

    int2 stack[kTraceStack];
    int sp = 0;
    int ref = m.rootRef;
    const NodePair* __restrict__ nodes = m.nodes;
    const TriBlock* __restrict__ tris = m.tris;

    for (;;) {
        if (ref >= 0) {
            const float4 a = __ldg(&nodes[ref].a), b = __ldg(&nodes[ref].b), c = __ldg(&nodes[ref].c);
            const int4 k = __ldg(&nodes[ref].d);
            if (STATS) st->pairs++;
            float ln, rn;
            bool hl, hr;
            slab_pair(rp, a, b, c, hl, ln, hr, rn);
            hl = hl && !(ln > best.t);
            hr = hr && !(rn > best.t);
            if (hl && hr) {
                int nearRef = k.x, farRef = k.y;
                float farT = rn;
                if (rn < ln) { nearRef = k.y; farRef = k.x; farT = ln; }
                if (sp < kTraceStack) stack[sp++] = make_int2(farRef, __float_as_int(farT));
                ref = nearRef;
                continue;
            }
            if (hl) { ref = k.x; continue; }
            if (hr) { ref = k.y; continue; }
        } else {
            const int packed = ~ref;
            const int first = packed >> kLeafCountBits;
            const int end = first + (packed & ((1 << kLeafCountBits) - 1)) + 1;
            for (int i = first; i < end; ++i) {
                const float4 p = __ldg(&tris[i].p), q = __ldg(&tris[i].q), r = __ldg(&tris[i].r);
                if (STATS) st->tris++;
                if (MODE != TRACE_ANY) {
                    const float facing = dot(v3(p.w, q.w, r.w), d);
                    if (MODE == TRACE_FORWARD && facing >= 0.0f) continue;
                    if (MODE == TRACE_BACKWARD && facing <= 0.0f) continue;
                }
                // intersectTriangleIndexed, reference src/mesh_intersection.cuh:15-46
                const V3 e1 = v3(q.x, q.y, q.z), e2 = v3(r.x, r.y, r.z);
                const V3 pv = cross(d, e2);
                const float det = dot(e1, pv);
                if (fabsf(det) < 1e-8f) continue;
                const float invDet = 1.0f / det;
                const V3 tv = o - v3(p.x, p.y, p.z);
                const float u = dot(tv, pv) * invDet;
                if (u < 0.0f || u > 1.0f) continue;
                const V3 qv = cross(tv, e1);
                const float v = dot(d, qv) * invDet;
                if (v < 0.0f || u + v > 1.0f) continue;
                const float t = dot(e2, qv) * invDet;
                if (t <= 1e-8f) continue;
                if (t > 1e-6f && t < best.t) best = Hit{t, u, v, static_cast<uint32_t>(i)};
            }
        }
        // pop: skip entries whose box entry distance is now beyond the best hit
        bool found = false;
        while (sp > 0) {
            const int2 e = stack[--sp];
            if (__int_as_float(e.y) > best.t) continue;
            ref = e.x;
            found = true;
            break;
        }
        if (!found) break;
    }
    return best;
}

// Out-of-line traversal with the mesh and the facing mode chosen at run time, per lane.  One copy of
// the loop per kernel instead of one per call site keeps the multi-trace stages (bounce entry,
// segment exits) inside the instruction cache, and lets lanes that are at different steps of a
// stage's trace sequence (or tracing different meshes) execute the same instructions together.
struct TraceTarget {
    const NodePair* nodes;
    const TriBlock* tris;
    int rootRef, numTriangles;
    V3 bmin, bmax;
};

__device__ __forceinline__ TraceTarget target_of(const MeshDev& m) {
    return TraceTarget{m.nodes, m.tris, m.rootRef, m.numTriangles, v3(m.rootMin[0], m.rootMin[1], m.rootMin[2]),
                       v3(m.rootMax[0], m.rootMax[1], m.rootMax[2])};
}

__device__ __forceinline__ TraceTarget pick_target(bool first, const TraceTarget& a, const TraceTarget& b) {
    TraceTarget t;
    t.nodes = first ? a.nodes : b.nodes;
    t.tris = first ? a.tris : b.tris;
    t.rootRef = first ? a.rootRef : b.rootRef;
    t.numTriangles = first ? a.numTriangles : b.numTriangles;
    t.bmin = first ? a.bmin : b.bmin;
    t.bmax = first ? a.bmax : b.bmax;
    return t;
}

// Control flow: "while-while" (Aila & Laine 2009).  Each lane descends inner nodes until it holds a
// leaf (or is done); the warp reconverges behind that loop, tests the leaves' triangles together,
// pops, and repeats.  Structured loops without `continue` give the compiler one reconvergence point
// per phase; the first version's single loop with early `continue`s left incoherent bounce rays at
// 7.5 of 32 active lanes (ncu, profiles/r1_v2_ncu_k_bounce_begin.txt).  Per-lane results and visit
// order are unchanged.
constexpr int kTraceDone = static_cast<int>(0x80000000u);  // not a leaf ref: first < 2^26 (checked at upload)
constexpr int kTracePop = static_cast<int>(0x80000001u);   // MultiTrace: the lane must pop its stack next

// tMax: only hits nearer than tMax are looked for; best.t stays tMax when there is none.
__device__ __forceinline__ Hit trace_dyn_inline(const TraceTarget& m, int mode, V3 o, V3 d, float tMax = 1e30f) {
    Hit best{tMax, 0.0f, 0.0f, 0u};
    if (m.numTriangles <= 0) return best;
    const V3 inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    float tn;
    if (!slab(o, inv, m.bmin.x, m.bmin.y, m.bmin.z, m.bmax.x, m.bmax.y, m.bmax.z, best.t, tn)) return best;
    const RayPack rp = ray_pack(o, inv);
    int2 stack[kTraceStack];
    int sp = 0;
    int ref = m.rootRef;
    const NodePair* __restrict__ nodes = m.nodes;
    const TriBlock* __restrict__ tris = m.tris;
    // facing modes skip subtrees whose normal cone says every triangle fails the facing test (nrb_scene.cuh)
    const int coneBit = mode == TRACE_FORWARD ? 1 : (mode == TRACE_BACKWARD ? 2 : 0);
    const float dlen = coneBit ? sqrtf(dot(d, d)) : 0.0f;
    // pop: skip entries whose box entry distance is now beyond the best hit
    auto pop = [&]() {
        int next = kTraceDone;
        while (sp > 0) {
            const int2 e = stack[--sp];
            if (!(__int_as_float(e.y) > best.t)) { next = e.x; break; }
        }
        return next;
    };
    while (ref != kTraceDone) {
        while (ref >= 0) {
            const float4 a = __ldg(&nodes[ref].a), b = __ldg(&nodes[ref].b), c = __ldg(&nodes[ref].c);
            const int4 k = __ldg(&nodes[ref].d);
            float ln, rn;
            bool hl, hr;
            slab_pair(rp, a, b, c, hl, ln, hr, rn);
            hl = hl && !(ln > best.t);
            hr = hr && !(rn > best.t);
            if (coneBit) {
                hl = hl && !(cone_culls(static_cast<uint32_t>(k.z), d.x, d.y, d.z, dlen) & coneBit);
                hr = hr && !(cone_culls(static_cast<uint32_t>(k.w), d.x, d.y, d.z, dlen) & coneBit);
            }
            if (hl && hr) {
                const bool rightFirst = rn < ln;
                if (sp < kTraceStack) stack[sp++] = make_int2(rightFirst ? k.x : k.y, __float_as_int(rightFirst ? ln : rn));
                ref = rightFirst ? k.y : k.x;
            } else if (hl) {
                ref = k.x;
            } else if (hr) {
                ref = k.y;
            } else {
                ref = pop();
            }
        }
        if (ref != kTraceDone) {
            const int packed = ~ref;
            const int first = packed >> kLeafCountBits;
            const int end = first + (packed & ((1 << kLeafCountBits) - 1)) + 1;
            for (int i = first; i < end; ++i) {
                const float4 p = __ldg(&tris[i].p), q = __ldg(&tris[i].q), r = __ldg(&tris[i].r);
                const float facing = dot(v3(p.w, q.w, r.w), d);
                bool ok = !((mode == TRACE_FORWARD && facing >= 0.0f) || (mode == TRACE_BACKWARD && facing <= 0.0f));
                const V3 e1 = v3(q.x, q.y, q.z), e2 = v3(r.x, r.y, r.z);
                const V3 pv = cross(d, e2);
                const float det = dot(e1, pv);
                ok = ok && !(fabsf(det) < 1e-8f);
                const float invDet = 1.0f / det;
                const V3 tv = o - v3(p.x, p.y, p.z);
                const float u = dot(tv, pv) * invDet;
                ok = ok && !(u < 0.0f || u > 1.0f);
                const V3 qv = cross(tv, e1);
                const float v = dot(d, qv) * invDet;
                ok = ok && !(v < 0.0f || u + v > 1.0f);
                const float t = dot(e2, qv) * invDet;
                ok = ok && !(t <= 1e-8f) && t > 1e-6f && t < best.t;
                if (ok) best = Hit{t, u, v, static_cast<uint32_t>(i)};
            }
            ref = pop();
        }
    }
    return best;
}

// ---- the same walk for a converged warp, reconverged every few steps ------------------------------------------------
// trace_dyn_inline leaves reconvergence to the compiler's placement: lanes of a warp drift apart over a long walk (one in
// its node loop, one testing a leaf, one popping) and each instruction issues for a handful of them.  trace_sync runs the
// identical per-lane walk in rounds of kSyncRound steps (a node pair or a leaf each) with a warp-wide vote between the
// rounds - a convergence barrier, nothing else: no state is exchanged, every lane visits what it would have visited.
// Measured on the ground-truth primary pass of the 1080p bench frame (k_gt_primary, ms): unbounded rounds 0.56 (= the
// loop above), 256 steps 0.56, 96: 0.46, 48: 0.43, 16: 0.39, 8: 0.39, 4: 0.377, 2: 0.379, 1: 0.381; with normal cones on
// the mesh (v24) 16: 0.349, 8: 0.345, 4: 0.326, 3: 0.311, 2: 0.318, 1: 0.305 - and k_gt_bounce 0.426 / 0.411 / 0.402 /
// 0.393 / 0.397 / 0.410: rounds of 2 give the shortest frame at 1 and at 8 samples per pixel.
// (Finishing the last long rays of a tile by all 32 lanes - each lane one slice of the ray's parameter range, exact by
// construction and parity-green - was also measured on top of the rounds: slower at every threshold, removed.)
// Every lane of the warp must call (valid = the lane holds a ray).
#ifndef NRB_SYNC_ROUND
#define NRB_SYNC_ROUND 2
#endif
constexpr int kSyncRound = NRB_SYNC_ROUND;

// TMIN: the caller knows that nothing it looks for lies nearer than tMin (the bounce entry's FORWARD search behind a
// nearest face that is a back face): subtrees the ray LEAVES before tMin (with a margin far above fp32 rounding) are
// skipped; hits are accepted as always, so the result is that of the unbounded walk.
template <bool STATS = false, bool TMIN = false>
__device__ __forceinline__ Hit trace_sync(const TraceTarget& m, int mode, bool valid, V3 o, V3 d, float tMax = 1e30f,
                                          TraceStats* st = nullptr, float tMin = -1e30f) {
    const unsigned full = 0xffffffffu;
    Hit best{tMax, 0.0f, 0.0f, 0u};
    int ref = kTraceDone;
    const V3 inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    if (valid && m.numTriangles > 0) {
        float tn;
        if (slab(o, inv, m.bmin.x, m.bmin.y, m.bmin.z, m.bmax.x, m.bmax.y, m.bmax.z, best.t, tn)) ref = m.rootRef;
    }
    const RayPack rp = ray_pack(o, inv);
    const float tMinCull = tMin - 1e-3f * (1.0f + fabsf(tMin));
    int2 stack[kTraceStack];
    int sp = 0;
    const NodePair* __restrict__ nodes = m.nodes;
    const TriBlock* __restrict__ tris = m.tris;
    // facing modes skip subtrees whose normal cone says every triangle fails the facing test (nrb_scene.cuh)
    const int coneBit = mode == TRACE_FORWARD ? 1 : (mode == TRACE_BACKWARD ? 2 : 0);
    const float dlen = coneBit ? sqrtf(dot(d, d)) : 0.0f;
    auto pop = [&]() {
        int next = kTraceDone;
        while (sp > 0) {
            const int2 e = stack[--sp];
            if (!(__int_as_float(e.y) > best.t)) { next = e.x; break; }
        }
        return next;
    };
    for (;;) {
        int steps = 0;
        while (ref != kTraceDone && steps < kSyncRound) {
            while (ref >= 0 && steps < kSyncRound) {
                ++steps;
                const float4 a = __ldg(&nodes[ref].a), b = __ldg(&nodes[ref].b), c = __ldg(&nodes[ref].c);
                const int4 k = __ldg(&nodes[ref].d);
                if (STATS) st->pairs++;
                float ln, rn;
                bool hl, hr;
                if (TMIN) {
                    float lf, rf;
                    slab_pair_far(rp, a, b, c, hl, ln, lf, hr, rn, rf);
                    hl = hl && !(lf < tMinCull);
                    hr = hr && !(rf < tMinCull);
                } else {
                    slab_pair(rp, a, b, c, hl, ln, hr, rn);
                }
                hl = hl && !(ln > best.t);
                hr = hr && !(rn > best.t);
                if (coneBit) {
                    hl = hl && !(cone_culls(static_cast<uint32_t>(k.z), d.x, d.y, d.z, dlen) & coneBit);
                    hr = hr && !(cone_culls(static_cast<uint32_t>(k.w), d.x, d.y, d.z, dlen) & coneBit);
                }
                if (hl && hr) {
                    const bool rightFirst = rn < ln;
                    if (sp < kTraceStack) stack[sp++] = make_int2(rightFirst ? k.x : k.y, __float_as_int(rightFirst ? ln : rn));
                    ref = rightFirst ? k.y : k.x;
                } else if (hl) {
                    ref = k.x;
                } else if (hr) {
                    ref = k.y;
                } else {
                    ref = pop();
                }
            }
            if (ref != kTraceDone && ref < 0) {
                ++steps;
                const int packed = ~ref;
                const int first = packed >> kLeafCountBits;
                const int end = first + (packed & ((1 << kLeafCountBits) - 1)) + 1;
                for (int i = first; i < end; ++i) {
                    const float4 p = __ldg(&tris[i].p), q = __ldg(&tris[i].q), r = __ldg(&tris[i].r);
                    if (STATS) st->tris++;
                    const float facing = dot(v3(p.w, q.w, r.w), d);
                    bool ok = !((mode == TRACE_FORWARD && facing >= 0.0f) || (mode == TRACE_BACKWARD && facing <= 0.0f));
                    const V3 e1 = v3(q.x, q.y, q.z), e2 = v3(r.x, r.y, r.z);
                    const V3 pv = cross(d, e2);
                    const float det = dot(e1, pv);
                    ok = ok && !(fabsf(det) < 1e-8f);
                    const float invDet = 1.0f / det;
                    const V3 tv = o - v3(p.x, p.y, p.z);
                    const float u = dot(tv, pv) * invDet;
                    ok = ok && !(u < 0.0f || u > 1.0f);
                    const V3 qv = cross(tv, e1);
                    const float v = dot(d, qv) * invDet;
                    ok = ok && !(v < 0.0f || u + v > 1.0f);
                    const float t = dot(e2, qv) * invDet;
                    ok = ok && !(t <= 1e-8f) && t > 1e-6f && t < best.t;
                    if (ok) best = Hit{t, u, v, static_cast<uint32_t>(i)};
                }
                ref = pop();
            }
        }
        if (!__any_sync(full, ref != kTraceDone)) break;
    }
    return best;
}

// The out-of-line copy.  A function that is really called keeps the global-memory descriptor in a vector
// register and moves it to a uniform one before every load (6 R2UR per node step): stages with one or two
// traces on their critical path (segment exits, re-entry) inline the loop instead, once per kernel.
static __device__ __noinline__ Hit trace_dyn(TraceTarget m, int mode, V3 o, V3 d) { return trace_dyn_inline(m, mode, o, d); }

// Slab test split into its best-independent part (ok) and the entry distance; the reference's full
// test is  ok && !(tNear > tMax).
__device__ __forceinline__ bool slab_base(V3 o, V3 inv, float mnx, float mny, float mnz, float mxx, float mxy,
                                          float mxz, float& tNear) {
    float t1 = (mnx - o.x) * inv.x, t2 = (mxx - o.x) * inv.x;
    float tmin = fminf(t1, t2), tmax = fmaxf(t1, t2);
    t1 = (mny - o.y) * inv.y; t2 = (mxy - o.y) * inv.y;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    t1 = (mnz - o.z) * inv.z; t2 = (mxz - o.z) * inv.z;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    tNear = tmin;
    return !(tmax < 0.0f || tmin > tmax);
}

// Three closest-hit queries along ONE ray in ONE depth-first pass: bit 0 = ANY, bit 1 = FORWARD_ONLY,
// bit 2 = BACKWARD_ONLY (`want` selects which are live).  Exact by construction: every node and
// stack entry carries the set of modes whose own traversal (traceMeshWithMode run separately for
// that mode) would have visited it - a mode stays in the set only while the node's entry distance
// is not beyond THAT mode's current best t - and the near/far order depends on the boxes only, so
// each mode sees its triangles in exactly the order of its separate traversal.  The bounce entry
// needs all three on the outer shell (early-termination ANY, entry FORWARD, exit BACKWARD when the
// origin is inside); one pass replaces up to three.
__device__ __noinline__ static void trace_multi(TraceTarget m, int want, V3 o, V3 d, float& tAny, float& tFwd,
                                                 float& tBwd) {
    float best0 = 1e30f, best1 = 1e30f, best2 = 1e30f;
    tAny = tFwd = tBwd = 1e30f;
    if (m.numTriangles <= 0 || want == 0) return;
    const V3 inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    float tn;
    if (!slab_base(o, inv, m.bmin.x, m.bmin.y, m.bmin.z, m.bmax.x, m.bmax.y, m.bmax.z, tn) || tn > 1e30f) return;
    const RayPack rp = ray_pack(o, inv);
    int2 stack[kTraceStack];
    unsigned char smask[kTraceStack];
    int sp = 0, ref = m.rootRef, mask = want;
    const NodePair* __restrict__ nodes = m.nodes;
    const TriBlock* __restrict__ tris = m.tris;
    auto live = [&](int set, float tNear) {
        int out = 0;
        if ((set & 1) && !(tNear > best0)) out |= 1;
        if ((set & 2) && !(tNear > best1)) out |= 2;
        if ((set & 4) && !(tNear > best2)) out |= 4;
        return out;
    };
    auto pop = [&]() {
        int next = kTraceDone;
        while (sp > 0) {
            --sp;
            const int again = live(smask[sp], __int_as_float(stack[sp].y));
            if (again) { next = stack[sp].x; mask = again; break; }
        }
        return next;
    };
    while (ref != kTraceDone) {  // while-while, see trace_dyn
        while (ref >= 0) {
            const float4 a = __ldg(&nodes[ref].a), b = __ldg(&nodes[ref].b), c = __ldg(&nodes[ref].c);
            const int4 k = __ldg(&nodes[ref].d);
            float ln, rn;
            bool bl, br;
            slab_pair(rp, a, b, c, bl, ln, br, rn);
            const int ml = bl ? live(mask, ln) : 0, mr = br ? live(mask, rn) : 0;
            if (ml && mr) {
                const bool rightFirst = rn < ln;
                if (sp < kTraceStack) {
                    stack[sp] = make_int2(rightFirst ? k.x : k.y, __float_as_int(rightFirst ? ln : rn));
                    smask[sp++] = static_cast<unsigned char>(rightFirst ? ml : mr);
                }
                ref = rightFirst ? k.y : k.x;
                mask = rightFirst ? mr : ml;
            } else if (ml) {
                ref = k.x;
                mask = ml;
            } else if (mr) {
                ref = k.y;
                mask = mr;
            } else {
                ref = pop();
            }
        }
        if (ref != kTraceDone) {
            const int packed = ~ref;
            const int first = packed >> kLeafCountBits;
            const int end = first + (packed & ((1 << kLeafCountBits) - 1)) + 1;
            for (int i = first; i < end; ++i) {
                const float4 p = __ldg(&tris[i].p), q = __ldg(&tris[i].q), r = __ldg(&tris[i].r);
                const float facing = dot(v3(p.w, q.w, r.w), d);
                int acc = mask & 1;
                if ((mask & 2) && !(facing >= 0.0f)) acc |= 2;
                if ((mask & 4) && !(facing <= 0.0f)) acc |= 4;
                const V3 e1 = v3(q.x, q.y, q.z), e2 = v3(r.x, r.y, r.z);
                const V3 pv = cross(d, e2);
                const float det = dot(e1, pv);
                bool ok = acc != 0 && !(fabsf(det) < 1e-8f);
                const float invDet = 1.0f / det;
                const V3 tv = o - v3(p.x, p.y, p.z);
                const float u = dot(tv, pv) * invDet;
                ok = ok && !(u < 0.0f || u > 1.0f);
                const V3 qv = cross(tv, e1);
                const float v = dot(d, qv) * invDet;
                ok = ok && !(v < 0.0f || u + v > 1.0f);
                const float t = dot(e2, qv) * invDet;
                ok = ok && !(t <= 1e-8f || !(t > 1e-6f));
                if (ok) {
                    if ((acc & 1) && t < best0) best0 = t;
                    if ((acc & 2) && t < best1) best1 = t;
                    if ((acc & 4) && t < best2) best2 = t;
                }
            }
            ref = pop();
        }
    }
    tAny = best0;
    tFwd = best1;
    tBwd = best2;
}

// The fused traversal as a resumable per-lane engine: the warp-level scheduler of a persistent kernel
// (k_bounce_begin) steps the lanes that hold a live trace and refills finished lanes with new work instead
// of letting them idle until the warp's longest ray ends.
//
// Pruning here is by the UNION of the live modes: a box is entered while its entry distance is not beyond
// `limit` = the largest best-t among the wanted modes (1e30 while one of them has no hit yet).
// (Round 2 experiment, removed: the nodes' normal cones - see trace_dyn_inline - as a second pruning rule for the
// FORWARD / BACKWARD members of the set.  Parity-green, but k_bounce_begin went from 0.543 to 0.624 ms: a bounce ray
// starts INSIDE the shell, where it lies inside the boxes of all coarse nodes - whose cones are too wide to cull -
// so the 36 node pairs per vertex are visited for every mode, and the cone decode only added instructions.)  trace_multi
// above tracks, per node, exactly which modes' own traversals would enter it (three compares per box, a
// mode set per stack entry); the union visits a superset of those nodes, and a mode m can gain a hit from
// the extra ones only through a triangle whose computed t is below best_m although its leaf box was entered
// beyond best_m - mathematically impossible (a hit inside a box is not nearer than the box), possible in
// fp32 only when t, best_m and the box distance agree to the last bits.  Entry distances are monotone down
// the tree (nested boxes, monotone fp32 slab arithmetic), so nothing else differs.  The parity hook
// NRB_TRACE_ALL_MODES_UNION runs this engine against three separate reference traversals.
struct MultiTrace {
    const NodePair* nodes;
    const TriBlock* tris;
    V3 d;
    RayPack rp;  // origin and reciprocal direction as fp32x2 couples (the origin is read back from it)
    float best0, best1, best2, limit;
    int ref, sp, want, cur;  // cur: next triangle of the current leaf, -1 = leaf not started
    int2 stack[kTraceStack];

    __device__ __forceinline__ bool done() const { return ref == kTraceDone; }
    __device__ __forceinline__ void relimit() {
        limit = fmaxf(fmaxf((want & 1) ? best0 : -1e30f, (want & 2) ? best1 : -1e30f), (want & 4) ? best2 : -1e30f);
    }
    // tMax: only hits nearer than tMax are looked for (best stays tMax when there is none)
    __device__ __forceinline__ void begin(const TraceTarget& m, int wanted, V3 org, V3 dir, float tMax = 1e30f) {
        best0 = best1 = best2 = tMax;
        limit = tMax;
        ref = kTraceDone;
        sp = 0;
        cur = -1;
        if (m.numTriangles <= 0 || wanted == 0) return;
        d = dir;
        const V3 inv = v3(1.0f / dir.x, 1.0f / dir.y, 1.0f / dir.z);
        float tn;
        if (!slab_base(org, inv, m.bmin.x, m.bmin.y, m.bmin.z, m.bmax.x, m.bmax.y, m.bmax.z, tn) || tn > tMax) return;
        rp = ray_pack(org, inv);
        nodes = m.nodes;
        tris = m.tris;
        want = wanted;
        ref = m.rootRef;
    }
    // Units of work for a warp-level scheduler that votes on what to run next: a node-pair step, one
    // triangle of the current leaf, or one stack entry of a pop.  kind(): 0 node, 1 triangle, 2 pop,
    // 3 finished.
    __device__ __forceinline__ int kind() const { return ref >= 0 ? 0 : (ref == kTraceDone ? 3 : (ref == kTracePop ? 2 : 1)); }
    __device__ __forceinline__ void step_node() {
        const float4 a = __ldg(&nodes[ref].a), b = __ldg(&nodes[ref].b), c = __ldg(&nodes[ref].c);
        const int4 k = __ldg(&nodes[ref].d);
        float ln, rn;
        bool hl, hr;
        slab_pair(rp, a, b, c, hl, ln, hr, rn);
        hl = hl && !(ln > limit);
        hr = hr && !(rn > limit);
        if (hl && hr) {
            const bool rightFirst = rn < ln;
            if (sp < kTraceStack) stack[sp++] = make_int2(rightFirst ? k.x : k.y, __float_as_int(rightFirst ? ln : rn));
            ref = rightFirst ? k.y : k.x;
        } else if (hl) {
            ref = k.x;
        } else if (hr) {
            ref = k.y;
        } else {
            ref = kTracePop;
        }
        cur = -1;
    }
    __device__ __forceinline__ void step_tri() {
        const int packed = ~ref;
        const int first = packed >> kLeafCountBits;
        const int end = first + (packed & ((1 << kLeafCountBits) - 1)) + 1;
        const int i = cur < 0 ? first : cur;
        const float4 p = __ldg(&tris[i].p), q = __ldg(&tris[i].q), r = __ldg(&tris[i].r);
        const float facing = dot(v3(p.w, q.w, r.w), d);
        int acc = want & 1;
        if ((want & 2) && !(facing >= 0.0f)) acc |= 2;
        if ((want & 4) && !(facing <= 0.0f)) acc |= 4;
        const V3 e1 = v3(q.x, q.y, q.z), e2 = v3(r.x, r.y, r.z);
        const V3 pv = cross(d, e2);
        const float det = dot(e1, pv);
        bool ok = acc != 0 && !(fabsf(det) < 1e-8f);
        const float invDet = 1.0f / det;
        V3 o;
        unpack2(rp.oxy, o.x, o.y);
        o.z = __uint_as_float(static_cast<unsigned int>(rp.ozz));
        const V3 tv = o - v3(p.x, p.y, p.z);
        const float u = dot(tv, pv) * invDet;
        ok = ok && !(u < 0.0f || u > 1.0f);
        const V3 qv = cross(tv, e1);
        const float v = dot(d, qv) * invDet;
        ok = ok && !(v < 0.0f || u + v > 1.0f);
        const float t = dot(e2, qv) * invDet;
        ok = ok && !(t <= 1e-8f || !(t > 1e-6f));
        if (ok) {
            if ((acc & 1) && t < best0) best0 = t;
            if ((acc & 2) && t < best1) best1 = t;
            if ((acc & 4) && t < best2) best2 = t;
            relimit();
        }
        cur = i + 1;
        if (cur >= end) ref = kTracePop;
    }
    __device__ __forceinline__ void step_pop() {
        if (sp == 0) {
            ref = kTraceDone;
        } else {
            --sp;
            if (!(__int_as_float(stack[sp].y) > limit)) {
                ref = stack[sp].x;
                cur = -1;
            }
        }
    }
};

template <bool STATS = false>
__device__ __forceinline__ Hit trace_mode(const MeshDev& m, int mode, V3 o, V3 d, TraceStats* st = nullptr) {
    if (mode == TRACE_FORWARD) return trace_closest<TRACE_FORWARD, STATS>(m, o, d, st);
    if (mode == TRACE_BACKWARD) return trace_closest<TRACE_BACKWARD, STATS>(m, o, d, st);
    return trace_closest<TRACE_ANY, STATS>(m, o, d, st);
}

// ---- hit record (computeHitData, reference src/mesh_intersection.cuh:168-238) -----------------

struct Surface {
    V3 position, normal;  // shading normal
    V2 uv;
    int materialId;
};

// sampleTextureRawDev, reference src/mesh_intersection.cuh:109-152 (8-bit texels, wrap, bilinear
// over (w-1) or nearest).  Returns x < 0 when the texture is unusable.
__device__ inline V3 tex_sample(const TexDev& t, float u, float v, bool nearest) {
    if (!t.pixels || t.width <= 0 || t.height <= 0 || t.channels <= 0) return v3(-1.0f, -1.0f, -1.0f);
    u = u - floorf(u);
    v = v - floorf(v);
    auto texel = [&](int xi, int yi) {
        const unsigned char* px = t.pixels + (yi * t.width + xi) * t.channels;
        const float k = 1.0f / 255.0f;
        return v3(px[0] * k, t.channels > 1 ? px[1] * k : 0.0f, t.channels > 2 ? px[2] * k : 0.0f);
    };
    if (nearest) {
        int x = static_cast<int>(u * static_cast<float>(t.width));
        int y = static_cast<int>(v * static_cast<float>(t.height));
        x = max(0, min(x, t.width - 1));
        y = max(0, min(y, t.height - 1));
        return texel(x, y);
    }
    const float x = u * static_cast<float>(t.width - 1), y = v * static_cast<float>(t.height - 1);
    const int x0 = static_cast<int>(floorf(x)), y0 = static_cast<int>(floorf(y));
    const int x1 = min(x0 + 1, t.width - 1), y1 = min(y0 + 1, t.height - 1);
    const float tx = x - static_cast<float>(x0), ty = y - static_cast<float>(y0);
    const V3 top = mix3(texel(x0, y0), texel(x1, y0), tx);
    const V3 bot = mix3(texel(x0, y1), texel(x1, y1), tx);
    return mix3(top, bot, ty);
}

// getMaterialIdForTriangle, reference src/mesh_intersection.cuh:53-85.
__device__ inline int material_of(const MeshDev& m, uint32_t tri) {
    if (!m.hasMeshMaterials || m.numPrimitives <= 0) return -1;
    uint32_t lo = 0, hi = static_cast<uint32_t>(m.numPrimitives);
    if (m.materialMap) {
        while (lo < hi) {
            const uint32_t mid = lo + (hi - lo) / 2;
            if (m.materialMap[mid] <= tri) lo = mid + 1; else hi = mid;
        }
    } else {
        lo = 0;
    }
    const uint32_t prim = lo > 0 ? lo - 1 : 0;
    if (prim < static_cast<uint32_t>(m.numPrimitives) && m.materialIds) return m.materialIds[prim];
    return -1;
}

__device__ inline Surface surface_at(const MeshDev& m, const Hit& h) {
    Surface s;
    const uint32_t i0 = m.indices[3 * h.prim], i1 = m.indices[3 * h.prim + 1], i2 = m.indices[3 * h.prim + 2];
    const float w = 1.0f - h.u - h.v;
    const TriBlock& tb = m.tris[h.prim];
    const V3 geo = v3(tb.p.w, tb.q.w, tb.r.w);
    s.normal = m.normals ? normalize(ld3(m.normals, i0) * w + ld3(m.normals, i1) * h.u + ld3(m.normals, i2) * h.v) : geo;
    s.position = ld3(m.vertices, i0) * w + ld3(m.vertices, i1) * h.u + ld3(m.vertices, i2) * h.v;
    if (m.texcoords) {
        const float* t = m.texcoords;
        s.uv.x = t[2 * i0] * w + t[2 * i1] * h.u + t[2 * i2] * h.v;
        s.uv.y = t[2 * i0 + 1] * w + t[2 * i1 + 1] * h.u + t[2 * i2 + 1] * h.v;
    } else {
        s.uv = V2{h.u, h.v};
    }
    s.materialId = material_of(m, h.prim);
    if (s.materialId >= 0 && s.materialId < m.numMaterials && m.materials) {
        const MaterialDev& mat = m.materials[s.materialId];
        if (mat.normal.textured && mat.normal.tex < static_cast<uint32_t>(m.numTextures)) {
            const V3 tn = tex_sample(m.textures[mat.normal.tex], s.uv.x, s.uv.y, m.textureNearest != 0);
            if (tn.x >= 0.0f) {
                const V3 ts = normalize(tn * 2.0f - v3s(1.0f));
                // orthoBasis, reference src/mesh_intersection.cuh:91-103
                V3 bt = v3s(0.0f);
                if (fabsf(s.normal.x) < 0.6f) bt.x = 1.0f;
                else if (fabsf(s.normal.y) < 0.6f) bt.y = 1.0f;
                else bt.z = 1.0f;
                const V3 tg = normalize(cross(bt, s.normal));
                bt = normalize(cross(s.normal, tg));
                s.normal = normalize(tg * ts.x + bt * ts.y + s.normal * ts.z);
            }
        }
    }
    return s;
}

// resolveMaterial restricted to what the kernels consume (albedo, metallic, roughness, specular):
// reference src/mesh_intersection.cuh:244-337 and the block at cuda_renderer_neural.cu:218-229.
__device__ inline float mat_scalar(const MatParam& p, V2 uv, const MeshDev& m) {
    if (!p.textured) return p.value;
    if (p.tex >= static_cast<uint32_t>(m.numTextures) || !m.textures) return p.value;
    const V3 s = tex_sample(m.textures[p.tex], uv.x, uv.y, m.textureNearest != 0);
    if (s.x < 0.0f) return p.value;
    return p.channel == 1 ? s.y : (p.channel == 2 ? s.z : s.x);
}

__device__ inline void resolve_surface_material(const MeshDev& m, const Surface& s, const MaterialDev& global,
                                                V3& albedo, V3& mparams) {
    const MaterialDev* mat = &global;
    if (s.materialId >= 0 && s.materialId < m.numMaterials && m.materials) mat = &m.materials[s.materialId];
    albedo = v3(mat->base_color.value[0], mat->base_color.value[1], mat->base_color.value[2]);
    if (mat->base_color.textured && mat->base_color.tex < static_cast<uint32_t>(m.numTextures) && m.textures) {
        const V3 raw = tex_sample(m.textures[mat->base_color.tex], s.uv.x, s.uv.y, m.textureNearest != 0);
        if (raw.x >= 0.0f) albedo = raw;  // (the reference's sRGB->linear step is commented out)
    }
    mparams = v3(mat_scalar(mat->metallic, s.uv, m), mat_scalar(mat->roughness, s.uv, m),
                 mat_scalar(mat->specular, s.uv, m));
}

}  // namespace nrb
