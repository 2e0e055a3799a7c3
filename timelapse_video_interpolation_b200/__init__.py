"""tvi — B200-native (sm_100a) kernels for the SVD keyframe-interpolation hot path.

Public surface (mirrors the reference's interfaces for this path):
  TviAttnProcessor, install, uninstall   diffusers AttentionProcessor plugin
  slerp                                  first/last-frame latent slerp
  cosine_similarity, backward_warp_check optical-flow checks
  feed_forward, layer_norm, install_feed_forward   FeedForward (GEGLU) / LayerNorm of the blocks
  install_blocks                         BasicTransformerBlock / TemporalBasicTransformerBlock forwards (N1, N2)
  install_spatio_temporal                TransformerSpatioTemporalModel.forward (GroupNorm, proj_in/out, AlphaBlender, layout)
  invalidate_caches                      drop the derived weight caches after `.data` surgery on the weights (EMA copy_to / restore)
  EulerCfgStepper, graphs.GraphedForward per-step glue (CFG + Euler + next model input in one launch), CUDA-graph capture (N4)
Everything runs through libtvi.so (include/tvi.h); there is no CPU or torch fallback.
"""
from . import _lib, ops  # noqa: F401
from ._lib import TviError, version  # noqa: F401
from .blocks import install_blocks, uninstall_blocks  # noqa: F401
from .caches import invalidate_caches  # noqa: F401
from .feedforward import feed_forward, install_feed_forward, layer_norm, uninstall_feed_forward  # noqa: F401
from .flow import backward_warp_check, check_clip_flows, compute_optical_flow_farneback, cosine_similarity, cosine_similarity_chw  # noqa: F401
from .processor import TviAttnProcessor, install, uninstall  # noqa: F401
from .slerp import slerp  # noqa: F401
from .step import EulerCfgStepper  # noqa: F401
from .spatio_temporal import install_spatio_temporal, uninstall_spatio_temporal  # noqa: F401

__all__ = ["TviAttnProcessor", "install", "uninstall", "slerp", "cosine_similarity", "cosine_similarity_chw",
           "backward_warp_check", "compute_optical_flow_farneback", "check_clip_flows", "feed_forward", "layer_norm", "install_feed_forward", "uninstall_feed_forward",
           "install_blocks", "uninstall_blocks", "EulerCfgStepper", "install_spatio_temporal", "uninstall_spatio_temporal",
           "TviError", "version", "ops", "invalidate_caches"]
