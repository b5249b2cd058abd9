"""Python mirror of the reference's operator interface for the hot path.

Same names and argument meaning as ``/root/reference/lightglue_pytorch_with_plugin/lightglue.py``:

* ``MHAHeadDim64`` (``lightglue.py:16-46``): ``symbolic`` emits the identical ONNX node
  ``CustomOp::MHAHeadDim64(query, key, value)`` (no attributes, output typed like ``query``), so an
  ONNX exported with this class is the graph the TensorRT plugin ``MHAHeadDim64`` / version ``1``
  is looked up for; ``forward`` runs the sm_100a kernel instead of eager SDPA.
* ``Attention`` (``lightglue.py:102-115``): module whose forward is ``MHAHeadDim64.apply``.

``bind_reference_model`` swaps these into an already-imported reference ``lightglue`` module so the
reference's own LightGlue (the caller of the path, 36 call sites per pair) drives the new kernel.
"""
from __future__ import annotations

import torch
from torch import nn

from .ops import attention

PLUGIN_NAME = "MHAHeadDim64"      # lightglue_attention_plugin.cpp:19
PLUGIN_VERSION = "1"              # lightglue_attention_plugin.cpp:20
PLUGIN_NAMESPACE = ""
ONNX_OP = "CustomOp::MHAHeadDim64"  # lightglue.py:22


class MHAHeadDim64(torch.autograd.Function):
    @staticmethod
    def symbolic(g, query, key, value):
        from torch.onnx import symbolic_helper as sh

        return g.op(ONNX_OP, query, key, value).setType(
            query.type().with_sizes([sh._get_tensor_dim_size(query, i) for i in range(4)])
        )

    @staticmethod
    def forward(ctx, query, key, value):
        # inference-only op, like the plugin (no backward in the reference either)
        return attention(query, key, value)


class Attention(nn.Module):
    def forward(self, query, key, value) -> torch.Tensor:
        return MHAHeadDim64.apply(query, key, value)


def bind_reference_model(lightglue_module) -> None:
    """Point an imported reference ``lightglue`` module's attention at the sm_100a kernel."""
    lightglue_module.MHAHeadDim64.forward = staticmethod(MHAHeadDim64.forward)
