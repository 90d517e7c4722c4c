"""hkm_assign_reduce_noalloc -- the Lloyd pass for embedders that may not allocate GPU memory
(a plugin living inside someone else's memory pool, e.g. a Spark executor or a serving arena).
It must give the same answer as hkm_assign_reduce, whose scratch the caller allocates, and it must
keep doing so on back-to-back calls even though its built-in scratch still holds the previous
call's partial sums.  Called straight through the C ABI with ctypes.

Collected last on purpose (file name): if this added entry point breaks, every other test has
already run."""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

VP, I32, I64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64


def _bind(hkm):
    lib = hkm._native.load()
    lib.hkm_assign_reduce_noalloc.argtypes = [VP, I64, I32, VP, I32, VP, VP, VP, VP, VP, I32, VP, I32, VP]
    lib.hkm_assign_reduce_noalloc.restype = I32
    return lib


def _ptr(t):
    return VP(t.data_ptr()) if t is not None else None


def _run_both(hkm, lib, n, d, k, seed, weighted):
    N = hkm._native
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    X = torch.randn(n, d, dtype=torch.float64, device="cuda", generator=g) * 1.5 + torch.randint(0, k, (n, 1), device="cuda", generator=g).double()
    C = X[torch.randperm(n, device="cuda", generator=g)[:k]].clone()
    w = (torch.rand(n, dtype=torch.float64, device="cuda", generator=g) * 2 + 0.1) if weighted else None
    P = k * d + k + 1
    stream = VP(torch.cuda.current_stream().cuda_stream)

    ref_partial = torch.empty(P, dtype=torch.float64, device="cuda")
    ref_labels = torch.empty(n, dtype=torch.int32, device="cuda"); ref_mind2 = torch.empty(n, dtype=torch.float64, device="cuda")
    scratch = N.new_scratch(d, k, "cuda")                      # the existing entry point: scratch from the caller
    N.assign_reduce(X, C, ref_partial, ref_labels, ref_mind2, weights=w, scratch=scratch)

    partial = torch.full((P,), float("nan"), dtype=torch.float64, device="cuda")
    labels = torch.full((n,), -7, dtype=torch.int32, device="cuda"); mind2 = torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
    rc = lib.hkm_assign_reduce_noalloc(_ptr(X), n, d, _ptr(C), k, _ptr(w), _ptr(partial), _ptr(labels), _ptr(mind2),
                                       None, 0, None, 1, stream)
    assert rc == 0, lib.hkm_last_error().decode()
    torch.cuda.synchronize()
    # same kernel, same grid, fixed summation order: the two entry points agree exactly, which is
    # tighter than the 1e-6 relative bound the parity tests allow
    assert torch.equal(labels, ref_labels)
    assert torch.equal(mind2, ref_mind2)
    assert torch.equal(partial, ref_partial)
    assert np.allclose(partial.cpu().numpy(), ref_partial.cpu().numpy(), rtol=1e-6)


def test_noalloc_matches_allocating_entry_point_twice_in_a_row(hkm):
    lib = _bind(hkm)
    # two calls in a row on different data and shapes: the second finds the first one's partial sums
    # (and a used ticket header) in the shared built-in scratch and must not be affected by them
    _run_both(hkm, lib, n=200_003, d=32, k=8, seed=1, weighted=False)
    _run_both(hkm, lib, n=50_000, d=16, k=16, seed=2, weighted=True)
    _run_both(hkm, lib, n=200_003, d=32, k=8, seed=3, weighted=False)


def test_noalloc_rejects_shapes_beyond_its_scratch(hkm):
    lib = _bind(hkm)
    n, d, k = 1000, 64, 64                                       # needs ~9.8 MB of scratch on a 148-SM part, the array holds 4 MiB
    X = torch.zeros(n, d, dtype=torch.float64, device="cuda"); C = torch.zeros(k, d, dtype=torch.float64, device="cuda")
    partial = torch.zeros(k * d + k + 1, dtype=torch.float64, device="cuda")
    assert hkm._native.assign_reduce_scratch_bytes(d, k) > (4 << 20)
    rc = lib.hkm_assign_reduce_noalloc(_ptr(X), n, d, _ptr(C), k, None, _ptr(partial), None, None, None, 0, None, 1,
                                       VP(torch.cuda.current_stream().cuda_stream))
    assert rc == -2                                              # HKM_ERR_UNSUPPORTED
    msg = lib.hkm_last_error().decode()
    assert "needs" in msg and "built-in array holds" in msg and "hkm_assign_reduce" in msg
