"""oracle/loader.py — TEST INFRASTRUCTURE.  Loads the CPU oracle libraries.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this.  'ref' = oracle/_ref/libnrb_ref.so (the reference's own headers compiled for the host, built
where /root/reference exists and shipped as a binary); 'port' = oracle/libnrb_oracle.so (our
restatement, built from source anywhere).
"""
import ctypes as C
import importlib
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_LIB = os.path.join(HERE, "libnrb_oracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libnrb_ref.so")
REFBVH_LIB = os.path.join(HERE, "_ref", "libnrb_refbvh.so")

_pkg = importlib.import_module("neural-renderer_b200")
_libs = {}


def have(kind):
    return os.path.exists(REF_LIB if kind == "ref" else PORT_LIB)


def oracle_lib(kind="auto"):
    if kind == "auto":
        kind = "ref" if have("ref") else "port"
    if kind not in _libs:
        _libs[kind] = _pkg.Lib(REF_LIB if kind == "ref" else PORT_LIB, "orc_")
    return _libs[kind]


def ref_build_bvh(vertices, indices):
    """Mesh::buildBvh of the reference (src/mesh_bvh.cpp:27-162) -> (nodes, reordered indices, bmin, bmax)."""
    dll = C.CDLL(REFBVH_LIB)
    v = np.ascontiguousarray(vertices, np.float32)
    idx = np.ascontiguousarray(indices, np.uint32).copy()
    nodes = np.zeros(2 * len(idx), _pkg.BVH_NODE_DTYPE)
    n = C.c_int(0)
    bmin = (C.c_float * 3)()
    bmax = (C.c_float * 3)()
    dll.ref_build_bvh(v.ctypes.data_as(C.c_void_p), C.c_int(len(v)), idx.ctypes.data_as(C.c_void_p),
                      C.c_int(len(idx)), nodes.ctypes.data_as(C.c_void_p), C.byref(n), bmin, bmax)
    return nodes[: n.value].copy(), idx, np.array(bmin, np.float32), np.array(bmax, np.float32)


def ref_extbvh_trace(vertices, indices, origins, directions, threads=0, reps=1, tmin=0.0):
    """ext/bvh's own CPU traversal (bvh::v2::Bvh::intersect + PrecomputedTri, the configuration of
    ext/bvh/test/benchmark.cpp) -> dict(prim = source triangle id or 0xffffffff, t, uv, normal, build_ms, trace_ms, threads)."""
    dll = C.CDLL(REFBVH_LIB)
    v = np.ascontiguousarray(vertices, np.float32)
    idx = np.ascontiguousarray(indices, np.uint32)
    o = np.ascontiguousarray(origins, np.float32)
    d = np.ascontiguousarray(directions, np.float32)
    n = len(o)
    prim, t = np.zeros(n, np.uint32), np.zeros(n, np.float32)
    uv, nrm = np.zeros((n, 2), np.float32), np.zeros((n, 3), np.float32)
    bms, tms, used = C.c_double(0), C.c_double(0), C.c_int(0)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    dll.ref_extbvh_trace(vp(v), C.c_int(len(v)), vp(idx), C.c_int(len(idx)), vp(o), vp(d), C.c_size_t(n), C.c_int(threads),
                         C.c_int(reps), C.c_float(tmin), vp(prim), vp(t), vp(uv), vp(nrm), C.byref(bms), C.byref(tms), C.byref(used))
    return dict(prim=prim, t=t, uv=uv, normal=nrm, build_ms=bms.value, trace_ms=tms.value, threads=used.value)


def have_refbvh():
    return os.path.exists(REFBVH_LIB)
