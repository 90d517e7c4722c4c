"""ctypes binding of csrc/libhkm.so (the C ABI in include/hkm.h).

There is deliberately NO fallback: if the CUDA library is missing or no GPU is present the
product path raises.  PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "csrc", "libhkm.so")

EXPORTS = (
    "hkm_last_error", "hkm_version", "hkm_max_k", "hkm_max_k_single_pass", "hkm_max_d", "hkm_needs_workspace", "hkm_set_lloyd_variant",
    "hkm_assign_reduce_scratch_bytes", "hkm_assign_reduce", "hkm_update_centers", "hkm_update_centers_cosine", "hkm_lloyd_step", "hkm_lloyd_step_peer", "hkm_peer_exchange_bytes", "hkm_peer_exchange_slot_doubles",
    "hkm_kmpar_scratch_bytes", "hkm_kmpar_cost_update", "hkm_kmpar_sample", "hkm_kmpar_strip_rows", "hkm_closest_counts",
    "hkm_silhouette_stats", "hkm_silhouette_score", "hkm_label_range", "hkm_label_counts",
    "hkm_assemble_scratch_bytes", "hkm_assemble", "hkm_assemble_filtered", "hkm_fill_f64", "hkm_moments_scratch_bytes", "hkm_moments", "hkm_column_moments", "hkm_scale_rows", "hkm_normalize_rows", "hkm_local_kmeanspp",
)

_lib = None


class HkmError(RuntimeError):
    pass


class LloydState(ctypes.Structure):
    _fields_ = [("cost", ctypes.c_double), ("iter", ctypes.c_int32), ("converged", ctypes.c_int32),
                ("max_shift2", ctypes.c_double), ("reserved", ctypes.c_int64)]


STATE_BYTES = ctypes.sizeof(LloydState)


def load() -> ctypes.CDLL:
    """Load libhkm.so; raises if it has not been built (python __graft_entry__.py / _build.py)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise HkmError(f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                           "There is no CPU fallback.")
        L = ctypes.CDLL(SO_PATH)
        vp, i32, i64, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_size_t
        L.hkm_last_error.restype = ctypes.c_char_p
        L.hkm_assign_reduce_scratch_bytes.restype = sz
        L.hkm_assign_reduce_scratch_bytes.argtypes = [i32, i32]
        L.hkm_assign_reduce.argtypes = [vp, i64, i32, vp, i32, vp, vp, vp, vp, vp, i32, vp, i32, vp, sz, vp]
        L.hkm_update_centers.argtypes = [vp, i32, i32, vp, ctypes.c_double, vp, vp]
        L.hkm_update_centers_cosine.argtypes = [vp, i32, i32, vp, ctypes.c_double, vp, vp]
        L.hkm_lloyd_step.argtypes = [vp, i64, i32, vp, i32, vp, vp, vp, i32, ctypes.c_double, vp, vp, sz, vp]
        L.hkm_peer_exchange_bytes.restype = sz
        L.hkm_peer_exchange_bytes.argtypes = [i32]
        L.hkm_lloyd_step_peer.argtypes = [vp, i64, i32, vp, i32, vp, vp, vp, i32, ctypes.c_double, vp, vp, sz, vp, i32, i32,
                                          ctypes.c_uint64, vp, vp]
        L.hkm_kmpar_scratch_bytes.restype = sz
        L.hkm_kmpar_scratch_bytes.argtypes = [i32]
        L.hkm_kmpar_cost_update.argtypes = [vp, i64, i32, vp, i32, vp, vp, vp, sz, vp]
        L.hkm_kmpar_sample.argtypes = [vp, i64, vp, vp, i64, i32, i32, ctypes.c_int32, i32, i32, vp, vp, i32, vp, vp]
        L.hkm_closest_counts.argtypes = [vp, i64, i32, vp, i32, vp, vp, vp, vp]
        L.hkm_assemble_scratch_bytes.restype = sz
        L.hkm_assemble_scratch_bytes.argtypes = [i64]
        L.hkm_assemble.argtypes = [vp, vp, vp, vp, i32, i64, vp, vp, i32, vp, vp, vp, vp, sz, vp]
        L.hkm_assemble_filtered.argtypes = [vp, vp, vp, vp, i32, i64, vp, vp, vp, vp, i64, i64, vp, i32, vp, vp, vp, vp, vp, vp, sz, vp]
        L.hkm_fill_f64.argtypes = [vp, i64, ctypes.c_double, vp]
        L.hkm_moments_scratch_bytes.restype = sz
        L.hkm_moments_scratch_bytes.argtypes = [i32]
        L.hkm_moments.argtypes = [vp, i64, i32, vp, vp, vp, sz, vp]
        L.hkm_column_moments.argtypes = [vp, vp, vp, vp, i32, i64, vp, vp, vp, sz, vp]
        L.hkm_scale_rows.argtypes = [vp, vp, i64, i32, vp, vp, vp]
        L.hkm_normalize_rows.argtypes = [vp, vp, i64, i32, vp, vp]
        L.hkm_local_kmeanspp.argtypes = [ctypes.c_int32, vp, vp, i32, i32, i32, i32, vp]
        L.hkm_silhouette_stats.argtypes = [vp, vp, vp, i64, i32, i32, vp, vp, sz, vp]
        L.hkm_label_range.argtypes = [vp, i64, vp, vp]
        L.hkm_label_counts.argtypes = [vp, i64, i32, vp, vp]
        L.hkm_silhouette_score.argtypes = [vp, vp, vp, i64, i32, i32, vp, vp, vp, vp, sz, vp]
        _lib = L
    return _lib


def _check(rc: int) -> None:
    if rc != 0:
        raise HkmError(f"libhkm error {rc}: {load().hkm_last_error().decode()}")


def require_cuda(t: torch.Tensor, name: str, dtype=None) -> None:
    if not t.is_cuda:
        raise HkmError(f"{name} must be a CUDA tensor (no CPU fallback)")
    if not t.is_contiguous():
        raise HkmError(f"{name} must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise HkmError(f"{name} must be {dtype}, got {t.dtype}")


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def set_lloyd_variant(variant: int) -> None:
    """hkm_set_lloyd_variant: pin one kernel family for the Lloyd passes (0 = automatic); tuning and tests only."""
    _check(load().hkm_set_lloyd_variant(int(variant)))


def assign_reduce_scratch_bytes(d: int, k: int) -> int:
    return int(load().hkm_assign_reduce_scratch_bytes(d, k))


def new_scratch(d: int, k: int, device) -> torch.Tensor:
    with torch.cuda.device(device):
        nbytes = assign_reduce_scratch_bytes(d, k)
    return torch.zeros((nbytes + 7) // 8, dtype=torch.int64, device=device)


def new_state(device) -> torch.Tensor:
    return torch.zeros(STATE_BYTES // 8, dtype=torch.int64, device=device)


def read_state(state: torch.Tensor) -> LloydState:
    raw = state.cpu().numpy().tobytes()
    return LloydState.from_buffer_copy(raw)


def assign_reduce(X, C, partial=None, labels=None, mind2=None, weights=None, state=None, scratch=None, do_reduce=True,
                  xnorm2_sum=None, xnorm2_mode=0):
    """One Lloyd data pass on the current stream (hkm_assign_reduce)."""
    require_cuda(X, "X", torch.float64)
    require_cuda(C, "C", torch.float64)
    n, d = X.shape
    k = C.shape[0]
    if load().hkm_needs_workspace(d, k):
        # more centres than one launch scores, or a shape outside the tiled kernels: the library runs the pass in chunks of
        # centres and needs per-row workspace
        if labels is None:
            labels = torch.empty(n, dtype=torch.int32, device=X.device)
        if mind2 is None:
            mind2 = torch.empty(n, dtype=torch.float64, device=X.device)
        xnorm2_sum, xnorm2_mode = None, 0
    with torch.cuda.device(X.device):
        rc = load().hkm_assign_reduce(_ptr(X), n, d, _ptr(C), k, _ptr(weights), _ptr(partial), _ptr(labels), _ptr(mind2),
                                      _ptr(xnorm2_sum), int(xnorm2_mode), _ptr(state), int(do_reduce), _ptr(scratch),
                                      scratch.numel() * scratch.element_size() if scratch is not None else 0, _stream())
    _check(rc)


def lloyd_step(X, C, partial, tol: float, state, scratch, weights=None, xnorm2_sum=None, xnorm2_mode=0):
    """hkm_lloyd_step: assign + reduce + centre update in one launch (single rank)."""
    require_cuda(X, "X", torch.float64)
    require_cuda(C, "C", torch.float64)
    n, d = X.shape
    k = C.shape[0]
    with torch.cuda.device(X.device):
        rc = load().hkm_lloyd_step(_ptr(X), n, d, _ptr(C), k, _ptr(weights), _ptr(partial), _ptr(xnorm2_sum), int(xnorm2_mode),
                                   float(tol), _ptr(state), _ptr(scratch), scratch.numel() * scratch.element_size(), _stream())
    _check(rc)


def lloyd_step_peer(X, C, partial, tol: float, state, scratch, xchg, weights=None, xnorm2_sum=None, xnorm2_mode=0):
    """hkm_lloyd_step_peer: one launch per Lloyd iteration on ranks > 1 (cross-GPU sum over NVLink peer memory)."""
    require_cuda(X, "X", torch.float64)
    require_cuda(C, "C", torch.float64)
    n, d = X.shape
    k = C.shape[0]
    xchg.epoch += 1
    with torch.cuda.device(X.device):
        rc = load().hkm_lloyd_step_peer(_ptr(X), n, d, _ptr(C), k, _ptr(weights), _ptr(partial), _ptr(xnorm2_sum), int(xnorm2_mode),
                                        float(tol), _ptr(state), _ptr(scratch), scratch.numel() * scratch.element_size(),
                                        ctypes.c_void_p(xchg.ptrs_dev), xchg.rank, xchg.world, xchg.epoch, _ptr(xchg.err), _stream())
    _check(rc)


def update_centers(partial, C, tol: float, state, cosine: bool = False):
    k, d = C.shape
    fn = load().hkm_update_centers_cosine if cosine else load().hkm_update_centers
    with torch.cuda.device(C.device):
        rc = fn(_ptr(partial), k, d, _ptr(C), float(tol), _ptr(state), _stream())
    _check(rc)


def kmpar_new_scratch(d: int, device) -> torch.Tensor:
    with torch.cuda.device(device):
        nbytes = int(load().hkm_kmpar_scratch_bytes(d))
    return torch.zeros((nbytes + 7) // 8, dtype=torch.int64, device=device)


def kmpar_cost_update(X, newC, costs, sum_out, scratch):
    """costs = min(costs, pointCost(newC, x)); sum_out[0] = costs.sum()  (hkm_kmpar_cost_update)."""
    require_cuda(X, "X", torch.float64); require_cuda(newC, "newC", torch.float64); require_cuda(costs, "costs", torch.float64)
    n, d = X.shape
    with torch.cuda.device(X.device):
        rc = load().hkm_kmpar_cost_update(_ptr(X), n, d, _ptr(newC), newC.shape[0], _ptr(costs), _ptr(sum_out), _ptr(scratch),
                                          scratch.numel() * scratch.element_size(), _stream())
    _check(rc)


def kmpar_strip_rows() -> int:
    return int(load().hkm_kmpar_strip_rows())


def kmpar_sample(costs, part_rows, strip_start, total_strips, part_index0, seed32, step, k, sum_costs, idx_out, count_out):
    require_cuda(costs, "costs", torch.float64)
    P = part_rows.numel() - 1
    with torch.cuda.device(costs.device):
        rc = load().hkm_kmpar_sample(_ptr(costs), costs.numel(), _ptr(part_rows), _ptr(strip_start), int(total_strips), P,
                                     int(part_index0), int(seed32), int(step), int(k), _ptr(sum_costs), _ptr(idx_out),
                                     idx_out.numel(), _ptr(count_out), _stream())
    _check(rc)


def closest_counts(X, cand, labels_work, mind2_work, counts_out):
    require_cuda(X, "X", torch.float64); require_cuda(cand, "cand", torch.float64)
    n, d = X.shape
    with torch.cuda.device(X.device):
        rc = load().hkm_closest_counts(_ptr(X), n, d, _ptr(cand), cand.shape[0], _ptr(labels_work), _ptr(mind2_work),
                                       _ptr(counts_out), _stream())
    _check(rc)


def label_counts(labels, m: int) -> torch.Tensor:
    """hkm_label_counts: occurrences of each label in [0, m) as an int64 tensor on the labels' device."""
    require_cuda(labels, "labels", torch.int32)
    counts = torch.empty(m, dtype=torch.int64, device=labels.device)
    _check(load().hkm_label_counts(_ptr(labels), labels.numel(), int(m), _ptr(counts), _stream()))
    return counts


def label_range(labels):
    """(min, max) of an int32 device column in one pass (hkm_label_range); (None, None) for an empty column."""
    require_cuda(labels, "labels", torch.int32)
    out = torch.empty(2, dtype=torch.int32, device=labels.device)
    with torch.cuda.device(labels.device):
        _check(load().hkm_label_range(_ptr(labels), labels.numel(), _ptr(out), _stream()))
    lo, hi = out.cpu().tolist()
    return (None, None) if labels.numel() == 0 else (lo, hi)


def silhouette_stats(X, labels, k, stats_out, scratch, weights=None):
    require_cuda(X, "X", torch.float64); require_cuda(labels, "labels", torch.int32)
    n, d = X.shape
    with torch.cuda.device(X.device):
        rc = load().hkm_silhouette_stats(_ptr(X), _ptr(labels), _ptr(weights), n, d, k, _ptr(stats_out), _ptr(scratch),
                                         scratch.numel() * scratch.element_size(), _stream())
    _check(rc)


def silhouette_score(X, labels, k, stats, work, out2, scratch, weights=None):
    require_cuda(X, "X", torch.float64); require_cuda(labels, "labels", torch.int32)
    n, d = X.shape
    with torch.cuda.device(X.device):
        rc = load().hkm_silhouette_score(_ptr(X), _ptr(labels), _ptr(weights), n, d, k, _ptr(stats), _ptr(work), _ptr(out2),
                                         _ptr(scratch), scratch.numel() * scratch.element_size(), _stream())
    _check(rc)


DTYPE_CODES = {torch.float64: 0, torch.float32: 1, torch.int32: 2, torch.int64: 3, torch.int16: 4, torch.int8: 5, torch.uint8: 6, torch.bool: 6}


def _col_tables(cols, valids):
    """HOST arrays of DEVICE pointers / dtype codes / validity kinds for the loader entry points."""
    d = len(cols)
    PtrArr, IntArr = ctypes.c_void_p * d, ctypes.c_int32 * d
    ptrs, dts, vptrs, vk = PtrArr(), IntArr(), PtrArr(), IntArr()
    for i, (c, v) in enumerate(zip(cols, valids)):
        require_cuda(c, f"column {i}")
        if c.dtype not in DTYPE_CODES:
            raise HkmError(f"column {i}: unsupported dtype {c.dtype}")
        ptrs[i], dts[i] = c.data_ptr(), DTYPE_CODES[c.dtype]
        if v is not None:
            require_cuda(v, f"validity {i}")
            vptrs[i], vk[i] = v.data_ptr(), 1      # HKM_VALID_BYTES
        else:
            vptrs[i], vk[i] = None, 0
    return ptrs, dts, vptrs, vk


def assemble(cols, valids, n, shift=None, scale=None, skip_invalid=False):
    """hkm_assemble: typed device columns -> row-major fp64 X (optionally standardised, optionally
    compacted).  Returns (X, n_out, invalid_rows)."""
    d = len(cols)
    dev = cols[0].device
    ptrs, dts, vptrs, vk = _col_tables(cols, valids)
    X = torch.empty((n, d), dtype=torch.float64, device=dev)
    n_out = torch.zeros(1, dtype=torch.int64, device=dev)
    invalid = torch.zeros(1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        sb = int(load().hkm_assemble_scratch_bytes(n)) if skip_invalid else 0
        scratch = torch.empty((sb + 7) // 8, dtype=torch.int64, device=dev) if sb else None
        rc = load().hkm_assemble(ptrs, dts, vptrs, vk, d, n, _ptr(shift), _ptr(scale), int(skip_invalid), _ptr(X), _ptr(n_out),
                                 _ptr(invalid), _ptr(scratch), sb, _stream())
    _check(rc)
    if not skip_invalid:
        return X, n, invalid          # no host sync: every row is kept; `invalid` is a device counter the caller may read
    n_keep, bad = int(n_out.item()), int(invalid.item())
    return X[:n_keep], n_keep, bad


def assemble_filtered(cols, valids, n, window=None, extra_valid=(), shift=None, scale=None):
    """hkm_assemble_filtered: WHERE ts BETWEEN lo AND hi + na.drop + VectorAssembler (+ scaling) in one compaction.
    window = (ts int64 device tensor, validity byte tensor or None, lo, hi) or None; extra_valid = validity byte tensors of
    columns that are not features.  Returns (X, keep mask (uint8 device tensor), rows outside the window, null rows inside it)."""
    d = len(cols)
    dev = cols[0].device
    ptrs, dts, vptrs, vk = _col_tables(cols, valids)
    X = torch.empty((n, d), dtype=torch.float64, device=dev)
    n_out = torch.zeros(1, dtype=torch.int64, device=dev)
    counts = torch.zeros(2, dtype=torch.int64, device=dev)
    keep = torch.zeros(n, dtype=torch.uint8, device=dev)
    ev = (ctypes.c_void_p * max(len(extra_valid), 1))()
    for i, v in enumerate(extra_valid):
        require_cuda(v, f"extra validity {i}", torch.uint8)
        ev[i] = v.data_ptr()
    ts, tsv, lo, hi = window if window is not None else (None, None, 0, 0)
    if ts is not None:
        require_cuda(ts, "window column", torch.int64)
    with torch.cuda.device(dev):
        sb = int(load().hkm_assemble_scratch_bytes(n))
        scratch = torch.empty((sb + 7) // 8, dtype=torch.int64, device=dev)
        rc = load().hkm_assemble_filtered(ptrs, dts, vptrs, vk, d, n, _ptr(shift), _ptr(scale), _ptr(ts), _ptr(tsv), int(lo), int(hi), ev, len(extra_valid),
                                          _ptr(X), _ptr(n_out), ctypes.c_void_p(counts.data_ptr()), ctypes.c_void_p(counts.data_ptr() + 8), _ptr(keep),
                                          _ptr(scratch), sb, _stream())
    _check(rc)
    n_keep = int(n_out.item())
    bad, outside = counts.cpu().tolist()
    return X[:n_keep], keep, outside, bad


def fill_f64(t, value: float) -> None:
    """hkm_fill_f64: t[:] = value on the current stream."""
    require_cuda(t, "tensor", torch.float64)
    with torch.cuda.device(t.device):
        _check(load().hkm_fill_f64(_ptr(t), t.numel(), float(value), _stream()))


def count_invalid_rows(cols, valids, n) -> int:
    """Rows holding a null or NaN (hkm_assemble in count-only mode: no matrix is written)."""
    d = len(cols)
    dev = cols[0].device
    ptrs, dts, vptrs, vk = _col_tables(cols, valids)
    n_out = torch.zeros(1, dtype=torch.int64, device=dev)
    invalid = torch.zeros(1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        rc = load().hkm_assemble(ptrs, dts, vptrs, vk, d, n, None, None, 0, None, _ptr(n_out), _ptr(invalid), None, 0, _stream())
    _check(rc)
    return int(invalid.item())


def moments(X, pivot):
    """hkm_moments -> tensor [d, 2] of (sum (x-K), sum (x-K)^2)."""
    require_cuda(X, "X", torch.float64)
    n, d = X.shape
    out = torch.zeros(2 * d, dtype=torch.float64, device=X.device)
    with torch.cuda.device(X.device):
        sb = int(load().hkm_moments_scratch_bytes(d))
        scratch = torch.empty((sb + 7) // 8, dtype=torch.int64, device=X.device)
        rc = load().hkm_moments(_ptr(X), n, d, _ptr(pivot), _ptr(out), _ptr(scratch), sb, _stream())
    _check(rc)
    return out.view(d, 2)


def column_moments(cols, valids, n, pivot):
    d = len(cols)
    dev = cols[0].device
    ptrs, dts, vptrs, vk = _col_tables(cols, valids)
    out = torch.zeros(2 * d, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        sb = int(load().hkm_moments_scratch_bytes(d))
        scratch = torch.empty((sb + 7) // 8, dtype=torch.int64, device=dev)
        rc = load().hkm_column_moments(ptrs, dts, vptrs, vk, d, n, _ptr(pivot), _ptr(out), _ptr(scratch), sb, _stream())
    _check(rc)
    return out.view(d, 2)


def scale_rows(X, shift, scale):
    require_cuda(X, "X", torch.float64)
    out = torch.empty_like(X)
    with torch.cuda.device(X.device):
        rc = load().hkm_scale_rows(_ptr(X), _ptr(out), X.shape[0], X.shape[1], _ptr(shift), _ptr(scale), _stream())
    _check(rc)
    return out


def normalize_rows(X):
    """hkm_normalize_rows -> (X_hat, number of zero-length rows)."""
    require_cuda(X, "X", torch.float64)
    out = torch.empty_like(X)
    zero = torch.zeros(1, dtype=torch.int64, device=X.device)
    with torch.cuda.device(X.device):
        rc = load().hkm_normalize_rows(_ptr(X), _ptr(out), X.shape[0], X.shape[1], _ptr(zero), _stream())
    _check(rc)
    return out, int(zero.item())


def local_kmeanspp(seed: int, points, weights, k: int, max_iter: int = 30):
    """hkm_local_kmeanspp (host arrays in, host array out): LocalKMeans.kMeansPlusPlus on the k-means|| candidates."""
    import numpy as np
    pts = np.ascontiguousarray(points, dtype=np.float64)
    wts = np.ascontiguousarray(weights, dtype=np.float64)
    m, d = pts.shape
    out = np.empty((k, d), dtype=np.float64)
    _check(load().hkm_local_kmeanspp(int(seed), pts.ctypes.data, wts.ctypes.data, m, d, int(k), int(max_iter), out.ctypes.data))
    return out
