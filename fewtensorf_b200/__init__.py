"""fewtensorf_b200 — B200-native (sm_100a) implementation of Few-TensoRF's per-ray hot path.

Public surface mirrors the reference: `TensorVMSplit`, `TensorCP`, `AlphaGridMask`
(models/tensoRF.py, models/tensorBase.py) and `OctreeRender_trilinear_fast` (renderer.py:13-26).
Importing this package loads libfewtensorf_b200.so and fails loudly if it is missing.
"""
from ._abi import LIB_PATH, lib  # noqa: F401  (import = load; raises if the library is absent)
from .module import (AlphaGridMask, MLPRender, MLPRender_Fea, MLPRender_PE, SHRender, TensorBase, TensorCP,  # noqa: F401
                     TensorVMSplit)
from .optim import FusedAdam, TVLoss  # noqa: F401
from .rays import DeviceRaySampler, generate_rays, render_view  # noqa: F401
from .renderer import OctreeRender_trilinear_fast  # noqa: F401

__all__ = ["TensorVMSplit", "TensorCP", "TensorBase", "AlphaGridMask", "OctreeRender_trilinear_fast",
           "MLPRender_Fea", "MLPRender_PE", "MLPRender", "SHRender", "FusedAdam", "TVLoss", "generate_rays", "render_view", "DeviceRaySampler",
           "LIB_PATH"]
