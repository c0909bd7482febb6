"""rtx_nerf_b200 — B200-native (sm_100a) implementation of owensgroup/rtx_nerf's per-ray hot path:
ray-grid intersection, sample generation, NeRF MLP evaluation and alpha compositing, behind the
reference's own sampler/, vol_render/ and rtx/ entry points (include/rtxnerf.h is the C ABI).

The CUDA library is mandatory: there is no CPU fallback anywhere in this package.
"""
from . import synth  # noqa: F401

__all__ = ["synth", "api", "build"]
