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
