"""ctypes binding of oracle/hkm_oracle.c plus a small NumPy twin.

TEST INFRASTRUCTURE ONLY: the product package never imports this module.  PARITY UNPINNED BY
THE REFERENCE (it holds no KMeans code or tests); pinned instead by the offline known answers
in tests/test_oracle_kat.py.  Each function cites the upstream Spark file it restates; the
reference itself only contributes the call idiom (ref:L135-136, L146-148, L162-163).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libhkm_oracle.so")


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "hkm_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        # -march=native code must be rebuilt on the box that runs it; fall back if unsupported.
        # -ffp-contract=off: the JVM never fuses a*b+c into an FMA, so the restatement must not either.
        cmd = ["gcc", "-O3", "-march=native", "-ffp-contract=off", "-fopenmp", "-fPIC", "-fvisibility=hidden", "-shared",
               "-o", _SO, src, "-lm"]
        subprocess.run(cmd, check=True, capture_output=True)
    return _SO


_lib = None
_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int32)
_lp = ctypes.POINTER(ctypes.c_int64)


def _built_for_this_cpu() -> bool:
    stamp = _SO + ".cpu"
    try:
        with open("/proc/cpuinfo") as f:
            model = next((l for l in f if l.startswith("model name")), "unknown").strip()
    except OSError:
        model = "unknown"
    ok = os.path.exists(stamp) and open(stamp).read() == model
    if not ok:
        os.makedirs(os.path.dirname(stamp), exist_ok=True)
        with open(stamp, "w") as f:
            f.write(model)
    return ok


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build(force=not _built_for_this_cpu())
        L = ctypes.CDLL(_SO)
        L.hkmo_hash_seed.restype = ctypes.c_uint64
        L.hkmo_hash_seed.argtypes = [ctypes.c_int64]
        for n in ("hkmo_xorshift_next_double", "hkmo_jrandom_next_double", "hkmo_xorshift_next_gaussian"):
            getattr(L, n).restype = ctypes.c_double
        for n in ("hkmo_xorshift_next_long", "hkmo_jrandom_next_long", "hkmo_take_sample"):
            getattr(L, n).restype = ctypes.c_int64
        L.hkmo_silhouette_sqeuclid.restype = ctypes.c_double
        L.hkmo_silhouette_cosine.restype = ctypes.c_double
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _c64(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a


class _State(ctypes.Structure):
    _fields_ = [("s", ctypes.c_uint64)]


class XORShiftRandom:
    """core/util/random/XORShiftRandom.scala (extends java.util.Random, overrides next)."""

    def __init__(self, seed: int):
        self._st = _State()
        self._have = ctypes.c_int(0)
        self._saved = ctypes.c_double(0.0)
        lib().hkmo_xorshift_init(ctypes.byref(self._st), ctypes.c_int64(_wrap64(seed)))

    def next_int(self) -> int:
        return lib().hkmo_xorshift_next_int(ctypes.byref(self._st))

    def next_long(self) -> int:
        return lib().hkmo_xorshift_next_long(ctypes.byref(self._st))

    def next_double(self) -> float:
        return lib().hkmo_xorshift_next_double(ctypes.byref(self._st))

    def next_gaussian(self) -> float:
        return lib().hkmo_xorshift_next_gaussian(ctypes.byref(self._st), ctypes.byref(self._have), ctypes.byref(self._saved))

    @property
    def state(self) -> int:
        return int(self._st.s)


class JavaRandom:
    """java.util.Random (48-bit LCG)."""

    def __init__(self, seed: int):
        self._st = _State()
        lib().hkmo_jrandom_init(ctypes.byref(self._st), ctypes.c_int64(_wrap64(seed)))

    def next_int(self, bound: int | None = None) -> int:
        if bound is None:
            return lib().hkmo_jrandom_next_int(ctypes.byref(self._st))
        return lib().hkmo_jrandom_next_int_bound(ctypes.byref(self._st), ctypes.c_int32(bound))

    def next_long(self) -> int:
        return lib().hkmo_jrandom_next_long(ctypes.byref(self._st))

    def next_double(self) -> float:
        return lib().hkmo_jrandom_next_double(ctypes.byref(self._st))


def _wrap64(v: int) -> int:
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >= (1 << 63) else v


def hash_seed(seed: int) -> int:
    return int(lib().hkmo_hash_seed(ctypes.c_int64(_wrap64(seed))))


def find_closest(C, x, use_stats=False, form=0):
    C = _c64(C); x = _c64(x)
    k, d = C.shape
    dist = ctypes.c_double()
    i = lib().hkmo_find_closest(_d(C), k, d, _d(x), int(use_stats), form, ctypes.byref(dist))
    return i, dist.value


def lloyd_pass(X, C, w=None, P=8, use_stats=False, form=0, threads=0, want_labels=False):
    """One mapPartitions + reduceByKey pass (mllib KMeans.runAlgorithmWithWeight body)."""
    X = _c64(X); C = _c64(C)
    n, d = X.shape; k = C.shape[0]
    sums = np.zeros((k, d)); wsum = np.zeros(k); cost = ctypes.c_double()
    labels = np.empty(n, dtype=np.int32) if want_labels else None
    wp = _d(_c64(w)) if w is not None else None
    lib().hkmo_lloyd_pass(_d(X), wp, ctypes.c_int64(n), d, _d(C), k, P, int(use_stats), form, threads,
                          _d(sums), _d(wsum), ctypes.byref(cost), labels.ctypes.data_as(_ip) if want_labels else None)
    return sums, wsum, cost.value, labels


def update_centers(sums, wsum, C, tol, form=0):
    C = np.array(C, dtype=np.float64, order="C")
    k, d = C.shape
    conv = lib().hkmo_update_centers(_d(_c64(sums)), _d(_c64(wsum)), k, d, ctypes.c_double(tol), form, _d(C))
    return C, bool(conv)


def lloyd(X, C0, w=None, max_iter=20, tol=1e-4, P=8, use_stats=False, form=0, threads=0):
    """Returns (centers, num_iter, training_cost)."""
    X = _c64(X); C = np.array(C0, dtype=np.float64, order="C")
    n, d = X.shape; k = C.shape[0]
    it = ctypes.c_int(); cost = ctypes.c_double()
    wp = _d(_c64(w)) if w is not None else None
    lib().hkmo_lloyd(_d(X), wp, ctypes.c_int64(n), d, _d(C), k, max_iter, ctypes.c_double(tol), P, int(use_stats), form, threads,
                     ctypes.byref(it), ctypes.byref(cost))
    return C, it.value, cost.value


def predict(X, C, use_stats=False, form=0, threads=0, want_dist=False):
    X = _c64(X); C = _c64(C)
    n, d = X.shape; k = C.shape[0]
    labels = np.empty(n, dtype=np.int32)
    dist = np.empty(n) if want_dist else None
    lib().hkmo_predict(_d(X), ctypes.c_int64(n), d, _d(C), k, int(use_stats), form, threads,
                       labels.ctypes.data_as(_ip), _d(dist) if want_dist else None)
    return (labels, dist) if want_dist else labels


def take_sample(n, P, num, seed):
    """RDD.takeSample(false, num, seed) over P contiguous partitions -> row indices."""
    out = np.empty(max(num, 1) if num < n else n, dtype=np.int64)
    m = lib().hkmo_take_sample(ctypes.c_int64(n), P, num, ctypes.c_int64(_wrap64(seed)), out.ctypes.data_as(_lp))
    return out[:m]


def init_random(X, k, seed, P=8):
    X = _c64(X); n, d = X.shape
    C = np.empty((min(k, n), d))
    m = lib().hkmo_init_random(_d(X), ctypes.c_int64(n), d, k, ctypes.c_int64(_wrap64(seed)), P, _d(C))
    return C[:m].copy()


def local_kmeanspp(seed, points, weights, k, max_iter=30, form=0):
    points = _c64(points); weights = _c64(weights)
    m, d = points.shape
    C = np.empty((k, d))
    lib().hkmo_local_kmeanspp(ctypes.c_int32(seed), _d(points), _d(weights), m, d, k, max_iter, form, _d(C))
    return C


def init_kmeans_parallel(X, k, seed, init_steps=2, P=8, first_idx=-1, form=0, threads=0, debug=False):
    X = _c64(X); n, d = X.shape
    C = np.empty((k, d))
    cap = 1 + 64 * k * init_steps + 1024
    cand = np.empty((cap, d)); ccount = ctypes.c_int()
    chosen = np.full(cap, -1, dtype=np.int64)
    steps = np.zeros(init_steps, dtype=np.int32)
    kout = lib().hkmo_init_kmeans_parallel(_d(X), ctypes.c_int64(n), d, k, init_steps, ctypes.c_int64(_wrap64(seed)), P,
                                           ctypes.c_int64(first_idx), form, threads, _d(C), _d(cand), cap, ctypes.byref(ccount),
                                           chosen.ctypes.data_as(_lp), steps.ctypes.data_as(_ip))
    if debug:
        tot = 1 + int(steps.sum())
        return C[:kout].copy(), dict(candidates=cand[:ccount.value].copy(), chosen_idx=chosen[:tot].copy(), step_counts=steps)
    return C[:kout].copy()


def silhouette(X, labels, w=None, kmax=None, P=8, threads=0, distance="squaredEuclidean"):
    X = _c64(X); labels = np.ascontiguousarray(labels, dtype=np.int32)
    n, d = X.shape
    kmax = int(labels.max()) + 1 if kmax is None else kmax
    wp = _d(_c64(w)) if w is not None else None
    if distance == "squaredEuclidean":
        return lib().hkmo_silhouette_sqeuclid(_d(X), labels.ctypes.data_as(_ip), wp, ctypes.c_int64(n), d, kmax, P, threads)
    return lib().hkmo_silhouette_cosine(_d(X), labels.ctypes.data_as(_ip), wp, ctypes.c_int64(n), d, kmax)


def scaler_fit(X, P=8):
    X = _c64(X); n, d = X.shape
    mean = np.empty(d); std = np.empty(d)
    lib().hkmo_scaler_fit(_d(X), ctypes.c_int64(n), d, P, _d(mean), _d(std))
    return mean, std


def scaler_transform(X, mean, std, with_mean=False, with_std=True):
    X = _c64(X); n, d = X.shape
    Y = np.empty_like(X)
    lib().hkmo_scaler_transform(_d(X), ctypes.c_int64(n), d, _d(_c64(mean)), _d(_c64(std)), int(with_mean), int(with_std), _d(Y))
    return Y


def num_threads() -> int:
    return lib().hkmo_num_threads()


FORM_COSINE = 2   # distanceMeasure="cosine" (CosineDistanceMeasure); 0 / 1 are the two Euclidean code paths


def kmeans_fit(X, k, seed, max_iter=20, tol=1e-4, init_mode="k-means||", init_steps=2, P=8, w=None,
               init_centers=None, first_idx=-1, use_stats=False, form=0, threads=0, distance="euclidean"):
    """ml.clustering.KMeans.fit restated end to end: init, Lloyd, summary pass.

    Returns dict(centers, num_iter, training_cost, labels, cluster_sizes)."""
    X = _c64(X)
    if distance == "cosine":
        form, use_stats = FORM_COSINE, False
    if init_centers is not None:
        C0 = np.array(init_centers, dtype=np.float64)
    elif init_mode == "random":
        C0 = init_random(X, k, seed, P)
    else:
        C0 = init_kmeans_parallel(X, k, seed, init_steps, P, first_idx, form, threads)
    C, it, cost = lloyd(X, C0, w, max_iter, tol, P, use_stats, form, threads)
    labels = predict(X, C, use_stats, form, threads)
    return dict(centers=C, num_iter=it, training_cost=cost, labels=labels,
                cluster_sizes=np.bincount(labels, minlength=C.shape[0]).astype(np.int64), init_centers=C0)


# ------------------------------------------------------------------------------------------
# NumPy twin (independent restatement used to cross-check the C code on small inputs)
# ------------------------------------------------------------------------------------------
def np_lloyd_pass(X, C, w=None):
    X = np.asarray(X, dtype=np.float64); C = np.asarray(C, dtype=np.float64)
    D = ((X[:, None, :] - C[None, :, :]) ** 2).sum(-1)
    lab = D.argmin(1)  # first minimum == strict '<' scan
    w = np.ones(len(X)) if w is None else np.asarray(w, dtype=np.float64)
    k = C.shape[0]
    sums = np.zeros_like(C); np.add.at(sums, lab, X * w[:, None])
    wsum = np.bincount(lab, weights=w, minlength=k)
    cost = float((D[np.arange(len(X)), lab] * w).sum())
    return sums, wsum, cost, lab.astype(np.int32)


def np_lloyd(X, C0, max_iter=20, tol=1e-4, w=None):
    C = np.array(C0, dtype=np.float64); it = 0; conv = False; cost = 0.0
    while it < max_iter and not conv:
        sums, wsum, cost, _ = np_lloyd_pass(X, C, w)
        conv = True
        for j in range(C.shape[0]):
            if wsum[j] > 0:
                nc = sums[j] * (1.0 / wsum[j])
                if ((nc - C[j]) ** 2).sum() > tol * tol:
                    conv = False
                C[j] = nc
        it += 1
    return C, it, cost


def np_silhouette(X, labels, w=None):
    X = np.asarray(X, dtype=np.float64); labels = np.asarray(labels)
    w = np.ones(len(X)) if w is None else np.asarray(w, dtype=np.float64)
    ids = np.unique(labels)
    sn = np.sqrt((X * X).sum(1)) ** 2
    Y = {c: (X[labels == c] * w[labels == c, None]).sum(0) for c in ids}
    Psi = {c: (sn[labels == c] * w[labels == c]).sum() for c in ids}
    N = {c: w[labels == c].sum() for c in ids}
    s = np.zeros(len(X))
    for i in range(len(X)):
        a = labels[i]
        if N[a] == w[i]:
            continue
        Dv = {c: sn[i] + Psi[c] / N[c] - 2 * (X[i] @ Y[c]) / N[c] for c in ids}
        cur = Dv[a] * N[a] / (N[a] - w[i]); nb = min(Dv[c] for c in ids if c != a)
        s[i] = 1 - cur / nb if cur < nb else (nb / cur - 1 if cur > nb else 0.0)
    return float((s * w).sum() / w.sum())
