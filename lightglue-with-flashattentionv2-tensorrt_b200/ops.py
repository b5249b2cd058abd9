"""Torch-tensor front end of the C-ABI launcher (device pointers + current stream only).

PyTorch is plumbing here: it owns device memory and the stream; the arithmetic happens in the
hand-written sm_100a kernel behind ``lg_attn_hd64_launch``.  There is no eager fallback — a CPU
tensor, a missing library or a non-sm_100 device raises.
"""
from __future__ import annotations

import torch

from . import _binding as B

HEAD_DIM = 64


def _mode(q: torch.Tensor, out_dtype: torch.dtype) -> int:
    if q.dtype == torch.float16 and out_dtype == torch.float16:
        return B.F16_F16
    if q.dtype == torch.float16 and out_dtype == torch.float32:
        return B.F16_F32
    if q.dtype == torch.float32 and out_dtype == torch.float32:
        return B.F32_F32
    raise TypeError(f"unsupported dtype combination: in {q.dtype}, out {out_dtype} "
                    "(supported: f16->f16, f16->f32, f32->f32)")


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, out_dtype: torch.dtype | None = None,
              out: torch.Tensor | None = None, workspace: torch.Tensor | None = None) -> torch.Tensor:
    """``softmax(q @ k^T / 8) @ v`` for ``q [B,H,N,64]``, ``k, v [B,H,M,64]`` on a B200.

    Semantics of the reference plugin call (``lightglue_attention_plugin.cpp:114-270``): all three
    inputs share a dtype (half or float), the output has ``q``'s shape; ``out_dtype`` additionally
    exposes the reference's internal half-in / float-out kernel (``…fp32out.cu:962-989``).
    """
    if not (q.is_cuda and k.is_cuda and v.is_cuda):
        raise RuntimeError("lg_attn: tensors must live on a CUDA device (no CPU fallback)")
    if q.dim() != 4 or k.dim() != 4 or v.dim() != 4:
        raise ValueError("lg_attn: q, k, v must be rank 4 [B,H,rows,64]")
    if q.shape[-1] != HEAD_DIM or k.shape[-1] != HEAD_DIM or v.shape[-1] != HEAD_DIM:
        raise ValueError("lg_attn: head_dim must be 64")
    if k.shape != v.shape or q.shape[:2] != k.shape[:2]:
        raise ValueError(f"lg_attn: shape mismatch q {tuple(q.shape)} k {tuple(k.shape)} v {tuple(v.shape)}")
    if not (q.dtype == k.dtype == v.dtype):
        raise TypeError("lg_attn: q, k, v must share a dtype")
    out_dtype = out_dtype or q.dtype
    mode = _mode(q, out_dtype)
    q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
    b, h, n, _ = q.shape
    m = k.shape[2]
    if out is None:
        out = torch.empty((b, h, n, HEAD_DIM), dtype=out_dtype, device=q.device)
    elif out.dtype != out_dtype or tuple(out.shape) != (b, h, n, HEAD_DIM) or not out.is_contiguous():
        raise ValueError("lg_attn: `out` has the wrong dtype/shape or is not contiguous")
    lib = B.load()
    ws_bytes = lib.lg_attn_hd64_workspace_bytes(b, h, n, m, mode)
    ws_ptr = None
    if ws_bytes:
        if workspace is None or workspace.numel() * workspace.element_size() < ws_bytes:
            workspace = torch.empty(ws_bytes, dtype=torch.uint8, device=q.device)
        ws_ptr = workspace.data_ptr()
    with torch.cuda.device(q.device):
        stream = torch.cuda.current_stream(q.device).cuda_stream
        rc = lib.lg_attn_hd64_launch(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(),
                                     b, h, n, m, mode, ws_ptr, ws_bytes, stream)
    B.check(rc, "lg_attn_hd64_launch")
    return out


def _bhr_strides(t: torch.Tensor) -> "B.Strides":
    return B.Strides(t.stride(0), t.stride(1), t.stride(2))


def attention_strided(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, out: torch.Tensor,
                      kv_batch_shift: int = 0) -> torch.Tensor:
    """Same attention on FP16 ``[B,H,rows,64]`` VIEWS with arbitrary batch / head / row strides
    (``lg_attn_hd64_launch_strided``): Q/K/V may point into packed projection outputs and ``out``
    (FP16 or FP32) into the buffer the next GEMM reads, so the head-split transposes and
    ``.contiguous()`` copies around every call in the reference (``lightglue.py:139-147,190-199``)
    disappear.  The last dimension must be contiguous; strides multiples of 16 bytes.
    ``kv_batch_shift``: batch entry ``b`` of ``q`` attends to entry ``(b + shift) % B`` of ``k`` / ``v``."""
    for t in (q, k, v, out):
        if not t.is_cuda:
            raise RuntimeError("lg_attn: tensors must live on a CUDA device (no CPU fallback)")
        if t.dim() != 4 or t.shape[-1] != HEAD_DIM or t.stride(-1) != 1:
            raise ValueError("lg_attn: expected [B,H,rows,64] views with a contiguous last dimension")
    if not (q.dtype == k.dtype == v.dtype == torch.float16) or out.dtype not in (torch.float16, torch.float32):
        raise TypeError("lg_attn: strided entry takes FP16 q,k,v and an FP16 or FP32 out")
    b, h, n, _ = q.shape
    m = k.shape[2]
    if k.shape != v.shape or tuple(k.shape[:2]) != (b, h) or tuple(out.shape) != (b, h, n, HEAD_DIM):
        raise ValueError("lg_attn: shape mismatch")
    sq, sk, sv, so = (_bhr_strides(t) for t in (q, k, v, out))
    with torch.cuda.device(q.device):
        stream = torch.cuda.current_stream(q.device).cuda_stream
        rc = B.load().lg_attn_hd64_launch_strided(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(),
                                                  b, h, n, m, int(out.dtype == torch.float32), sq, sk, sv, so,
                                                  int(kv_batch_shift), stream)
    B.check(rc, "lg_attn_hd64_launch_strided")
    return out


def attention_host(q, k, v, out_dtype=None, device: int = 0):
    """Host-buffer entry (numpy arrays in, numpy array out) through ``lg_attn_hd64_run_host``."""
    import numpy as np

    q = np.ascontiguousarray(q)
    k = np.ascontiguousarray(k)
    v = np.ascontiguousarray(v)
    out_dtype = np.dtype(out_dtype or q.dtype)
    key = (q.dtype.name, out_dtype.name)
    modes = {("float16", "float16"): B.F16_F16, ("float16", "float32"): B.F16_F32,
             ("float32", "float32"): B.F32_F32}
    if key not in modes:
        raise TypeError(f"unsupported dtype combination {key}")
    b, h, n, d = q.shape
    m = k.shape[2]
    if d != HEAD_DIM or k.shape != v.shape or k.shape[:2] != (b, h) or k.shape[3] != HEAD_DIM:
        raise ValueError("lg_attn: bad shapes")
    o = np.empty((b, h, n, HEAD_DIM), dtype=out_dtype)
    rc = B.load().lg_attn_hd64_run_host(q.ctypes.data, k.ctypes.data, v.ctypes.data, o.ctypes.data,
                                        b, h, n, m, modes[key], device)
    B.check(rc, "lg_attn_hd64_run_host")
    return o


class HostSession:
    """Host-buffer front end (``lg_attn_session_*``): persistent device buffers, round-robin streams.

    ``submit`` takes host arrays / pinned CPU tensors (anything with a stable address) and is
    asynchronous; results are valid after ``sync``.
    """

    def __init__(self, max_batch: int, heads: int, max_n_q: int, max_n_kv: int, mode: int = B.F16_F16,
                 n_slots: int = 4, device: int = 0):
        import ctypes

        self._lib = B.load()
        self._h = ctypes.c_void_p()
        self.heads = heads
        B.check(self._lib.lg_attn_session_create(ctypes.byref(self._h), device, max_batch, heads, max_n_q, max_n_kv,
                                                 mode, n_slots), "lg_attn_session_create")

    def submit(self, q_ptr: int, k_ptr: int, v_ptr: int, o_ptr: int, batch: int, n_q: int, n_kv: int) -> None:
        B.check(self._lib.lg_attn_session_submit(self._h, q_ptr, k_ptr, v_ptr, o_ptr, batch, n_q, n_kv),
                "lg_attn_session_submit")

    def submit_many(self, q_ptr: int, k_ptr: int, v_ptr: int, o_ptr: int, count: int, q_stride: int, kv_stride: int, out_stride: int,
                    batch: int, n_q: int, n_kv: int, group: int = 4) -> None:
        """``count`` same-shape problems; packed host arrays ([q | k | v] per problem, dense outputs) move ``group`` problems per
        PCIe transfer each way (``lg_attn_session_submit_many``)."""
        B.check(self._lib.lg_attn_session_submit_many(self._h, q_ptr, k_ptr, v_ptr, o_ptr, count, q_stride, kv_stride, out_stride, group, batch,
                                                      n_q, n_kv), "lg_attn_session_submit_many")

    def sync(self) -> None:
        B.check(self._lib.lg_attn_session_sync(self._h), "lg_attn_session_sync")

    def close(self) -> None:
        if self._h:
            self._lib.lg_attn_session_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
