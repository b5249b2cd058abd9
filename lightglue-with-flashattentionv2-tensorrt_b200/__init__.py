"""B200-native (sm_100a) drop-in for the fused-attention hot path of
qdLMF/LightGlue-with-FlashAttentionV2-TensorRT (plugin ``MHAHeadDim64``).

Only what the path needs lives here: ``csrc/`` (CUDA kernel, C-ABI, TensorRT plugin TU) and the
host-side mirror of the reference's operator interface.  See DESIGN.md / INTEGRATION.md.
"""
from . import _binding, matcher, sharding
from ._binding import C_ABI_SYMBOLS, F16_F16, F16_F32, F32_F32, LIB_NAME, LgAttnError, load
from .matcher import NativeMatcher, PairMatcher, reference_init_state_dict
from .ops import HEAD_DIM, HostSession, attention, attention_host, attention_strided
from .plugin_op import (ONNX_OP, PLUGIN_NAME, PLUGIN_NAMESPACE, PLUGIN_VERSION, Attention,
                        MHAHeadDim64, bind_reference_model)

__all__ = [
    "PairMatcher", "NativeMatcher", "reference_init_state_dict", "matcher", "sharding", "attention", "attention_host", "attention_strided", "HostSession", "Attention", "MHAHeadDim64", "bind_reference_model", "load",
    "LgAttnError", "HEAD_DIM", "F16_F16", "F16_F32", "F32_F32", "C_ABI_SYMBOLS", "LIB_NAME",
    "PLUGIN_NAME", "PLUGIN_VERSION", "PLUGIN_NAMESPACE", "ONNX_OP",
]
