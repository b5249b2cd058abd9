"""GPU: the TensorRT plugin boundary, driven by the mock host through TensorRT's call sequence
(registry -> createPlugin/deserializePlugin -> format negotiation -> configurePlugin with dynamic
dims -> getWorkspaceSize -> clone -> enqueue), for both I/O types and dynamic N/M (configs[1],[2])."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import attention_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def host(pkg):
    lib = pkg.load()
    lib.lg_mock_trt_enqueue.restype = ctypes.c_int
    lib.lg_mock_trt_enqueue.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int] * 3 + [ctypes.c_void_p, ctypes.c_size_t,
                                                                                   ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
    lib.lg_mock_trt_workspace_bytes.restype = ctypes.c_size_t
    lib.lg_mock_trt_workspace_bytes.argtypes = [ctypes.c_int]
    return lib


def _enqueue(host, q, k, v, is_float, via_deserialize=0, use_clone=1):
    n, m = q.shape[2], k.shape[2]
    ws_bytes = host.lg_mock_trt_workspace_bytes(is_float)
    ws = torch.empty(max(ws_bytes, 16), dtype=torch.uint8, device=DEV)   # TensorRT owns the workspace
    o = torch.empty(1, 4, n, 64, dtype=q.dtype, device=DEV)
    rc = host.lg_mock_trt_enqueue(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), n, m, is_float, ws.data_ptr(),
                                  ws_bytes, torch.cuda.current_stream().cuda_stream, via_deserialize, use_clone)
    assert rc == 0, rc
    torch.cuda.synchronize()
    return o


@pytest.mark.parametrize("n,m", [(1024, 1024), (1024, 1000), (512, 1024), (777, 1024), (1, 1), (2048, 2048), (1000, 37)])
def test_plugin_half(pkg, host, n, m):
    q, k, v = orc.make_qkv(n * 5 + m, 1, 4, n, m)
    tq, tk, tv = (torch.from_numpy(x).to(DEV) for x in (q, k, v))
    o = _enqueue(host, tq, tk, tv, 0)
    assert orc.max_abs_err(o.float().cpu().numpy(), orc.attention_np(q, k, v)) <= orc.TOL_F16_MAX_ABS
    assert torch.equal(o, pkg.attention(tq, tk, tv))       # same launcher underneath


@pytest.mark.parametrize("n,m", [(1024, 1024), (513, 1024), (640, 1024), (1023, 1024), (2048, 1)])
def test_plugin_float(pkg, host, n, m):
    q, k, v = orc.make_qkv(n * 3 + m, 1, 4, n, m, dtype=np.float32)
    tq, tk, tv = (torch.from_numpy(x).to(DEV) for x in (q, k, v))
    o = _enqueue(host, tq, tk, tv, 1)
    ref = orc.attention_np(q.astype(np.float16), k.astype(np.float16), v.astype(np.float16))
    assert o.dtype == torch.float32
    assert orc.rel_err(o.cpu().numpy(), ref) <= orc.TOL_F32_REL


def test_plugin_deserialized_and_unclone_paths(host):
    q, k, v = orc.make_qkv(1, 1, 4, 300, 200)
    tq, tk, tv = (torch.from_numpy(x).to(DEV) for x in (q, k, v))
    a = _enqueue(host, tq, tk, tv, 0, via_deserialize=0, use_clone=1)
    b = _enqueue(host, tq, tk, tv, 0, via_deserialize=1, use_clone=1)
    c = _enqueue(host, tq, tk, tv, 0, via_deserialize=1, use_clone=0)
    assert torch.equal(a, b) and torch.equal(a, c)


def test_plugin_dynamic_shapes_back_to_back(host):
    """One engine, changing keypoint counts call after call (trtexec profile min 1 / max 2048, steps.txt:152)."""
    rng = np.random.default_rng(0)
    for _ in range(12):
        n, m = int(rng.integers(1, 2049)), int(rng.integers(1, 2049))
        q, k, v = orc.make_qkv(n + m, 1, 4, n, m)
        tq, tk, tv = (torch.from_numpy(x).to(DEV) for x in (q, k, v))
        o = _enqueue(host, tq, tk, tv, 0)
        assert orc.max_abs_err(o.float().cpu().numpy(), orc.attention_np(q, k, v)) <= orc.TOL_F16_MAX_ABS, (n, m)


def test_python_operator_mirror_on_gpu(pkg):
    """`Attention()(q,k,v)` — the reference's module name and call shape (with_plugin/lightglue.py:102-115)."""
    q, k, v = orc.make_qkv(8, 1, 4, 1024, 1024)
    tq, tk, tv = (torch.from_numpy(x).to(DEV) for x in (q, k, v))
    o = pkg.Attention()(tq, tk, tv)
    sdpa = torch.nn.functional.scaled_dot_product_attention(tq.float(), tk.float(), tv.float())
    assert (o.float() - sdpa).abs().max().item() <= orc.TOL_F16_MAX_ABS
