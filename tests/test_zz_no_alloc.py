"""GPU: lg_match_assign_static — the assignment + mutual-NN filter with its scratch in a file-scope device array (no allocation in
the call; include/lg_match.h).  No counterpart in the reference.  Collected last on purpose.

The new entry point is called through the C-ABI (ctypes) and compared with the existing entry point, lg_match_assign with caller
scratch, on the same inputs: the kernels and their order are identical, so the outputs must be bit-equal (the tolerance the matcher
tests use for assign-vs-assign comparisons), on two calls in a row — the second sees the first call's leftovers in the scratch.
"""
import ctypes

import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    from lgattn_pkg import pkg

    return pkg, pkg._binding.load()


def _inputs(seed, b, n0, n1, dev):
    g = torch.Generator(device="cpu").manual_seed(seed)
    sim = (torch.randn(b, n0, n1, generator=g) * 3.0).to(dev).contiguous()
    z0 = torch.randn(b * n0, generator=g).half().to(dev).contiguous()
    z1 = torch.randn(b * n1, generator=g).half().to(dev).contiguous()
    return sim, z0, z1


def _run_existing(lib, sim, z0, z1, b, n0, n1, thr, with_scores):
    dev = sim.device
    m = torch.full((b, n0), -7, dtype=torch.int32, device=dev)
    ms = torch.full((b, n0), -7.0, dtype=torch.float32, device=dev)
    sc = torch.full((b, n0, n1), -7.0, dtype=torch.float32, device=dev) if with_scores else None
    nb = lib.lg_match_assign_workspace_bytes(b, n0, n1)
    ws = torch.empty(nb, dtype=torch.uint8, device=dev)
    rc = lib.lg_match_assign(sim.data_ptr(), z0.data_ptr(), z1.data_ptr(), 1, b, n0, n1, thr, m.data_ptr(), ms.data_ptr(),
                             sc.data_ptr() if with_scores else None, ws.data_ptr(), nb, None)
    assert rc == 0, rc
    torch.cuda.synchronize()
    return m, ms, sc


def _run_static(lib, sim, z0, z1, b, n0, n1, thr, outs):
    m, ms, sc = outs
    rc = lib.lg_match_assign_static(sim.data_ptr(), z0.data_ptr(), z1.data_ptr(), 1, b, n0, n1, thr, m.data_ptr(), ms.data_ptr(),
                                    sc.data_ptr() if sc is not None else None, None)
    assert rc == 0, rc
    torch.cuda.synchronize()


def test_static_scratch_entry_matches_existing_entry_twice_in_a_row(env):
    pkg, lib = env
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    thr = 0.0
    # two different shapes back to back (ragged, scalar path; then a vectorised shape with several row blocks): the second call
    # finds the first call's contents in the shared scratch
    cases = [(11, 2, 301, 263, True), (12, 3, 768, 1024, True), (13, 1, 64, 64, False)]
    refs = [_run_existing(lib, *_inputs(seed, b, n0, n1, dev), b, n0, n1, thr, ws) for (seed, b, n0, n1, ws) in cases]
    # outputs are allocated BEFORE the calls under test; the calls themselves must not allocate
    outs = []
    for (seed, b, n0, n1, ws) in cases:
        outs.append((torch.full((b, n0), -9, dtype=torch.int32, device=dev), torch.full((b, n0), -9.0, dtype=torch.float32, device=dev),
                     torch.full((b, n0, n1), -9.0, dtype=torch.float32, device=dev) if ws else None))
    ins = [_inputs(seed, b, n0, n1, dev) for (seed, b, n0, n1, ws) in cases]
    torch.cuda.synchronize()
    for rep in range(2):
        before = torch.cuda.memory_allocated(dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        for (case, inp, out) in zip(cases, ins, outs):
            _, b, n0, n1, _ = case
            _run_static(lib, *inp, b, n0, n1, thr, out)
        assert torch.cuda.memory_allocated(dev) == before
        if rep == 1:   # first pass may load the module lazily; after it the driver's free memory must not move either
            assert torch.cuda.mem_get_info(dev)[0] == free_before
        for (case, ref, out) in zip(cases, refs, outs):
            assert torch.equal(out[0], ref[0]), (rep, case)
            assert torch.equal(out[1], ref[1]), (rep, case)
            if ref[2] is not None:
                assert torch.equal(out[2], ref[2]), (rep, case)
            assert int((ref[0] >= 0).sum()) > 0   # the comparison is not vacuous: there are matches
        for out in outs:   # fresh sentinels for the second pass
            out[0].fill_(-9)
            out[1].fill_(-9.0)
            if out[2] is not None:
                out[2].fill_(-9.0)
        if rep == 0:
            torch.cuda.synchronize()


def test_static_scratch_entry_rejects_what_does_not_fit(env):
    pkg, lib = env
    dev = torch.device("cuda", 0)
    b, n0, n1 = 128, 1024, 1024
    assert lib.lg_match_assign_workspace_bytes(b, n0, n1) > pkg._binding.ASSIGN_STATIC_MAX_BYTES
    one = torch.zeros(16, dtype=torch.float32, device=dev)
    p = one.data_ptr()
    rc = lib.lg_match_assign_static(p, p, p, 1, b, n0, n1, 0.0, p, p, None, None)
    assert rc == pkg._binding.ESTATIC_TOO_LARGE
    assert "LG_MATCH_ASSIGN_STATIC_MAX_BYTES" in pkg._binding.error_string(rc)
    assert lib.lg_match_assign_static(p, p, p, 1, 0, 8, 8, 0.0, p, p, None, None) == pkg._binding.EINVAL
    torch.cuda.synchronize()
