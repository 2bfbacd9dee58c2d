"""ctypes binding of the C ABI declared in include/nrb200.h.

``Lib`` binds any library exporting these entry points under a prefix; the product is
``nrb_`` in libnrb200.so.  (The CPU oracle exports the same calls as ``orc_`` and is loaded only by
oracle/loader.py, for tests and the bench baseline.)  Struct layouts mirror the reference's host structs (src/bvh_data.h:5-13,
src/material.h:9-117, src/mesh.h:23-78).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


class BvhNode(C.Structure):
    _fields_ = [("bounds_min", C.c_float * 3), ("bounds_max", C.c_float * 3), ("left", C.c_int32),
                ("right", C.c_int32), ("first", C.c_int32), ("count", C.c_int32), ("is_leaf", C.c_int32)]


BVH_NODE_DTYPE = np.dtype([("bounds_min", "<f4", 3), ("bounds_max", "<f4", 3), ("left", "<i4"), ("right", "<i4"),
                           ("first", "<i4"), ("count", "<i4"), ("is_leaf", "<i4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim", "<u4")])


class MaterialParam(C.Structure):
    _fields_ = [("value", C.c_float), ("texture_id", C.c_uint32), ("channel", C.c_uint8),
                ("textured", C.c_uint8), ("pad_", C.c_uint8 * 2)]


class MaterialParam3(C.Structure):
    _fields_ = [("value", C.c_float * 3), ("texture_id", C.c_uint32), ("textured", C.c_uint8),
                ("pad_", C.c_uint8 * 3)]


class Material(C.Structure):
    _fields_ = [("base_color", MaterialParam3), ("metallic", MaterialParam), ("roughness", MaterialParam),
                ("specular", MaterialParam), ("specular_tint", MaterialParam), ("anisotropy", MaterialParam),
                ("sheen", MaterialParam), ("sheen_tint", MaterialParam), ("clearcoat", MaterialParam),
                ("clearcoat_gloss", MaterialParam), ("normal", MaterialParam3), ("base_emission", MaterialParam3),
                ("emission_scale", C.c_float), ("ior", C.c_float), ("specular_transmission", C.c_float)]

    @staticmethod
    def default() -> "Material":
        """Material::defaultMaterial(), reference src/material.h:86-106."""
        m = Material()
        m.base_color.value[:] = [0.8, 0.8, 0.8]
        m.clearcoat_gloss.value = 1.0
        m.emission_scale = 1.0
        m.ior = 1.5
        return m

    @staticmethod
    def from_config(base_color=(1.0, 1.0, 1.0), roughness=1.0, metallic=0.0, specular=0.0, specular_tint=0.0,
                    anisotropy=0.0, sheen=0.0, sheen_tint=0.0, clearcoat=0.0, clearcoat_gloss=0.0) -> "Material":
        """applyMaterialConfig of the evaluate driver (reference src/evaluate.cu:181-209) on the default."""
        m = Material.default()
        m.base_color.value[:] = list(base_color)
        for k, v in dict(roughness=roughness, metallic=metallic, specular=specular, specular_tint=specular_tint,
                         anisotropy=anisotropy, sheen=sheen, sheen_tint=sheen_tint, clearcoat=clearcoat,
                         clearcoat_gloss=clearcoat_gloss).items():
            getattr(m, k).value = v
        return m


class Texture(C.Structure):
    _fields_ = [("pixels", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32), ("channels", C.c_int32),
                ("srgb", C.c_int32)]


class MeshDesc(C.Structure):
    _fields_ = [("vertices", C.c_void_p), ("normals", C.c_void_p), ("texcoords", C.c_void_p),
                ("indices", C.c_void_p), ("num_triangles", C.c_int32), ("num_vertices", C.c_int32),
                ("nodes", C.c_void_p), ("num_nodes", C.c_int32), ("materials", C.c_void_p),
                ("material_map", C.c_void_p), ("material_ids", C.c_void_p), ("num_primitives", C.c_int32),
                ("num_materials", C.c_int32), ("textures", C.c_void_p), ("num_textures", C.c_int32),
                ("texture_nearest", C.c_int32), ("has_mesh_materials", C.c_int32), ("bounds_min", C.c_float * 3),
                ("bounds_max", C.c_float * 3)]


class NNConfig(C.Structure):
    _fields_ = [("log2_hashmap_size", C.c_int32), ("base_resolution", C.c_int32), ("n_points", C.c_int32),
                ("n_levels", C.c_int32), ("n_features_per_level", C.c_int32), ("per_level_scale", C.c_float),
                ("sh_degree", C.c_int32), ("mlp_width", C.c_int32), ("n_hidden_layers", C.c_int32)]


assert C.sizeof(BvhNode) == 44 and C.sizeof(Material) == 180 and C.sizeof(MaterialParam) == 12

SLOT_ORIGINAL, SLOT_INNER, SLOT_OUTER, SLOT_ADDITIONAL = 0, 1, 2, 3
TRACE_ANY, TRACE_FORWARD_ONLY, TRACE_BACKWARD_ONLY = 0, 1, 2
(OPT_USE_NEURAL_QUERY, OPT_SAMPLES_PER_PIXEL, OPT_BOUNCE_COUNT, OPT_LAMBERT_VIEW, OPT_CLASSIC_MESH_INDEX,
 OPT_USE_ADDITIONAL_MESH, OPT_USE_CONSTANT_NEURAL_COLOR, OPT_USE_DIRECT_ENV_COLOR, OPT_CAPTURE,
 OPT_USE_HARDWARE_RT, OPT_PROFILE, OPT_QUERY_KERNEL, OPT_TAIL_GRAPH, OPT_DENOISE, OPT_LITERAL_BOUNCE_TRACES, OPT_PATH_TAIL, OPT_TILE_ORDER, OPT_FRAME_SLICES,
 OPT_PIPELINE, OPT_BOUNCE_ENTRY) = range(20)

PRODUCT_LIB = os.path.join(HERE, "libnrb200.so")


class NrbError(RuntimeError):
    pass


def _vp(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Lib:
    """One loaded library + the prefix of its entry points."""

    def __init__(self, path: str, prefix: str):
        if not os.path.exists(path):
            raise NrbError(f"{path} is missing — run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no fallback implementation)")
        self.path, self.prefix = path, prefix
        self.dll = C.CDLL(path)
        self._sig()

    def fn(self, name):
        return getattr(self.dll, self.prefix + name)

    def has(self, name):
        return hasattr(self.dll, self.prefix + name)

    def _sig(self):
        vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
        P = C.POINTER
        table = {
            "create": ([P(NNConfig), i, P(vp)], i), "destroy": ([vp], None), "last_error": ([], C.c_char_p),
            "update_mesh": ([vp, i, P(MeshDesc)], i), "update_env": ([vp, vp, i, i, f], i),
            "set_global_material": ([vp, P(Material)], i), "load_weights": ([vp, vp, sz, i], i),
            "load_weights_file": ([vp, C.c_char_p, i], i), "params_bytes": ([vp], sz),
            "weights_file_bytes": ([vp], sz), "resize": ([vp, i, i], i),
            "set_camera_basis": ([vp, P(f), P(f), P(f), f], i), "set_option": ([vp, i, i], i),
            "get_option": ([vp, i], i), "set_envmap_rotation": ([vp, f], i),
            "set_constant_neural_color": ([vp, P(f)], i), "set_direct_env_color": ([vp, P(f)], i),
            "reset_samples": ([vp], i), "set_tile": ([vp, i, i, i], i), "render": ([vp, P(f)], i),
            "device_pixels": ([vp], vp), "read_pixels": ([vp, vp, sz], i), "read_pixels_async": ([vp, vp, sz], i), "width": ([vp], i), "height": ([vp], i),
            "accum_sample_count": ([vp], C.c_uint32), "get_timings": ([vp, P(C.c_double), P(i), P(C.c_uint64)], i),
            "get_counters": ([vp, P(C.c_uint64)], i), "get_trace_stats": ([vp, P(C.c_uint64)], i), "read_buffer": ([vp, C.c_char_p, vp, sz], i),
            "trace_rays": ([vp, i, i, vp, vp, sz, vp], i), "trace_rays_host": ([vp, i, i, vp, vp, sz, vp], i),
            "neural_query": ([vp, vp, sz, vp], i), "neural_query_host": ([vp, vp, sz, vp], i), "neural_query_static": ([vp, vp, sz, vp], i),
            "input_dims": ([vp], i), "output_dims": ([vp], i), "sync": ([vp], i), "stream": ([vp], vp),
            "gather_export": ([vp, i, vp, sz], i), "gather_attach": ([vp, vp, sz], i), "gather_detach": ([vp], i),
            "build_bvh": ([vp, i, vp, i, vp, P(i), vp], i),
            "build_bvh_gpu": ([vp, i, vp, i, vp, P(i), vp, i, i, P(C.c_float)], i), "build_bvh_gpu_error": ([], C.c_char_p), "abi_version": ([], i),
            "get_stage": ([vp, i, P(C.c_char_p), P(C.c_double), P(C.c_uint64)], i), "reset_stages": ([vp], i),
            "set_threads": ([vp, i], i), "set_row_range": ([vp, i, i], i), "set_row_sample": ([vp, i, i, i], i), "backend": ([], C.c_char_p),
        }
        for name, (args, res) in table.items():
            if self.has(name):
                fn = self.fn(name)
                fn.argtypes, fn.restype = args, res

    def check(self, code, what=""):
        if code != 0:
            msg = self.fn("last_error")()
            raise NrbError(f"{self.prefix}{what} failed ({code}): {msg.decode() if msg else ''}")


_libs: dict = {}


def product_lib() -> Lib:
    """libnrb200.so — fails loudly when it has not been built (no CPU fallback exists)."""
    if "nrb" not in _libs:
        _libs["nrb"] = Lib(PRODUCT_LIB, "nrb_")
    return _libs["nrb"]
