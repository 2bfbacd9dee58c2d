"""nrb200 — B200-native hot path of Alehandreus/neural-renderer.

The compute lives in libnrb200.so (hand-written sm_100a CUDA behind the C ABI of include/nrb200.h);
this package is the host-side mirror of the reference's renderer interface plus scene/weight
generators.  There is no CPU or PyTorch fallback: a missing libnrb200.so is an error.
"""
from ._capi import (BVH_NODE_DTYPE, HIT_DTYPE, Lib, Material, MeshDesc, NNConfig, NrbError, PRODUCT_LIB,
                    SLOT_ADDITIONAL, SLOT_INNER, SLOT_ORIGINAL, SLOT_OUTER, TRACE_ANY, TRACE_BACKWARD_ONLY,
                    TRACE_FORWARD_ONLY, product_lib)
from .renderer import MeshData, RenderBasis, RendererNeural, Scene, compute_camera_basis, look_at_basis
from . import evaluate, metrics, scenes, tiles

# torch_net (the PyTorch twin of the network + weight-file exporter) is imported on demand:
#   importlib.import_module("neural-renderer_b200.torch_net")
__all__ = ["RendererNeural", "Scene", "MeshData", "RenderBasis", "Material", "NNConfig", "scenes", "tiles", "evaluate", "metrics",
           "compute_camera_basis", "look_at_basis", "product_lib", "Lib", "NrbError"]
