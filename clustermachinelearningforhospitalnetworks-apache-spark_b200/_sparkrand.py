"""Host-side restatement of the small sequential pieces of Spark's KMeans initialisation.

These run on the driver in Spark too (they touch O(k) rows, not the data set):
  * XORShiftRandom + hashSeed        (core/util/random/XORShiftRandom.scala)      SURVEY 8a-6
  * java.util.Random                  (LocalKMeans, RDD.takeSample, randomizeInPlace)
  * RDD.takeSample(false, num, seed)  (core/rdd/RDD.scala + SamplingUtils + BernoulliSampler /
                                       GapSampling; O(#samples) thanks to gap sampling)  8a-6/8a-8
  * LocalKMeans.kMeansPlusPlus        (mllib/clustering/LocalKMeans.scala)         8a-7
Product code: written independently of oracle/ (which tests use to cross-check it).
"""
from __future__ import annotations

import math
from typing import List, Sequence

import numpy as np

_M64 = (1 << 64) - 1
_M32 = (1 << 32) - 1


def _i32(v: int) -> int:
    v &= _M32
    return v - (1 << 32) if v & (1 << 31) else v


def _i64(v: int) -> int:
    v &= _M64
    return v - (1 << 64) if v & (1 << 63) else v


def _rotl32(x: int, r: int) -> int:
    return ((x << r) | (x >> (32 - r))) & _M32


def _mm3_bytes_hash(data: bytes, seed: int) -> int:
    """scala.util.hashing.MurmurHash3.bytesHash."""
    h = seed & _M32
    n = len(data)
    i = 0
    while n - i >= 4:
        k = data[i] | (data[i + 1] << 8) | (data[i + 2] << 16) | (data[i + 3] << 24)
        k = (k * 0xCC9E2D51) & _M32
        k = _rotl32(k, 15)
        k = (k * 0x1B873593) & _M32
        h ^= k
        h = _rotl32(h, 13)
        h = (h * 5 + 0xE6546B64) & _M32
        i += 4
    rem = n - i
    k = 0
    if rem == 3:
        k ^= data[i + 2] << 16
    if rem >= 2:
        k ^= data[i + 1] << 8
    if rem >= 1:
        k ^= data[i]
        k = (k * 0xCC9E2D51) & _M32
        k = _rotl32(k, 15)
        k = (k * 0x1B873593) & _M32
        h ^= k
    h ^= n
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & _M32
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & _M32
    h ^= h >> 16
    return h


def hash_seed(seed: int) -> int:
    """XORShiftRandom.hashSeed: two MurmurHash3 rounds over the big-endian seed bytes."""
    b = (seed & _M64).to_bytes(8, "big")
    lo = _mm3_bytes_hash(b, 0x3C074A61)
    hi = _mm3_bytes_hash(b, lo)
    return (hi << 32) | lo


class XORShiftRandom:
    def __init__(self, seed: int):
        self.s = hash_seed(seed)

    def next(self, bits: int) -> int:
        x = self.s
        x ^= (x << 21) & _M64
        x ^= x >> 35
        x ^= (x << 4) & _M64
        self.s = x
        return _i32(x & ((1 << bits) - 1))

    def next_int(self) -> int:
        return self.next(32)

    def next_double(self) -> float:
        return ((self.next(26) << 27) + self.next(27)) * (1.0 / (1 << 53))


class JavaRandom:
    _MASK = (1 << 48) - 1

    def __init__(self, seed: int):
        self.s = ((seed & _M64) ^ 0x5DEECE66D) & self._MASK

    def next(self, bits: int) -> int:
        self.s = (self.s * 0x5DEECE66D + 0xB) & self._MASK
        return _i32(self.s >> (48 - bits))

    def next_int(self, bound: int | None = None) -> int:
        if bound is None:
            return self.next(32)
        r = self.next(31)
        m = bound - 1
        if bound & m == 0:
            return (bound * r) >> 31
        u = r
        while True:
            r = u % bound
            if _i32(u - r + m) >= 0:
                return r
            u = self.next(31)

    def next_long(self) -> int:
        return _i64((self.next(32) << 32) + self.next(32))

    def next_double(self) -> float:
        return ((self.next(26) << 27) + self.next(27)) * (1.0 / (1 << 53))


def partition_bounds(sizes: Sequence[int]) -> List[int]:
    out = [0]
    for s in sizes:
        out.append(out[-1] + int(s))
    return out


def equal_partition_sizes(n: int, P: int) -> List[int]:
    """ParallelCollectionRDD.slice positions: [floor(i*n/P), floor((i+1)*n/P))."""
    return [((i + 1) * n) // P - (i * n) // P for i in range(P)]


def _bernoulli_partition(lo: int, hi: int, fraction: float, pseed: int) -> List[int]:
    if fraction <= 0.0:
        return []
    if fraction >= 1.0:
        return list(range(lo, hi))
    rng = XORShiftRandom(pseed)
    out: List[int] = []
    if fraction <= 0.4:  # RandomSampler.defaultMaxGapSamplingFraction -> GapSampling
        lnq = math.log1p(-fraction)

        def gap() -> int:
            u = max(rng.next_double(), 5e-11)  # RandomSampler.rngEpsilon
            q = math.log(u) / lnq
            return 2147483647 if q >= 2147483647.0 else int(q)

        if hi > lo:  # lazy val gapSampling: first advance() happens on the first sample() call
            r = lo + gap()
            while r < hi:
                out.append(r)
                r += 1 + gap()
    else:
        for r in range(lo, hi):
            if rng.next_double() <= fraction:
                out.append(r)
    return out


def take_sample(part_sizes: Sequence[int], num: int, seed: int) -> List[int]:
    """RDD.takeSample(withReplacement=false, num, seed) -> global row indices (in Spark's order)."""
    bounds = partition_bounds(part_sizes)
    n = bounds[-1]
    if num == 0 or n == 0:
        return []
    rand = JavaRandom(seed)

    def shuffle(a: List[int]) -> List[int]:
        for i in range(len(a) - 1, 0, -1):
            j = rand.next_int(i + 1)
            a[j], a[i] = a[i], a[j]
        return a

    if num >= n:
        return shuffle(list(range(n)))
    fr0 = num / n
    gamma = -math.log(1e-4) / n
    fraction = min(1.0, max(1e-10, fr0 + gamma + math.sqrt(gamma * gamma + 2 * gamma * fr0)))
    while True:
        prng = JavaRandom(rand.next_int())
        samples: List[int] = []
        for p in range(len(part_sizes)):
            samples += _bernoulli_partition(bounds[p], bounds[p + 1], fraction, prng.next_long())
        if len(samples) >= num:
            break
    return shuffle(samples)[:num]


def distinct_rows(V: np.ndarray) -> np.ndarray:
    """Seq.distinct on dense vectors: first occurrence kept."""
    seen = set()
    keep = []
    for i, row in enumerate(V):
        key = row.tobytes() if not np.any(row == 0) else (row + 0.0).tobytes()  # -0.0 == 0.0
        if key not in seen:
            seen.add(key)
            keep.append(i)
    return V[keep]


def _find_closest(C: np.ndarray, x: np.ndarray) -> int:
    d2 = ((C - x) ** 2).sum(axis=1)
    return int(np.argmin(d2))  # first minimum == strict '<'


def local_kmeans_pp(seed: int, points: np.ndarray, weights: np.ndarray, k: int, max_iter: int = 30) -> np.ndarray:
    """LocalKMeans.kMeansPlusPlus: weighted k-means++ then <= max_iter Lloyd steps on the candidates."""
    rand = JavaRandom(seed)
    m, d = points.shape
    centers = np.empty((k, d))
    # pickWeighted
    wsum = 0.0
    for w in weights:
        wsum += float(w)
    r = rand.next_double() * wsum
    i, cur = 0, 0.0
    while i < m and cur < r:
        cur += float(weights[i])
        i += 1
    centers[0] = points[max(i - 1, 0)]
    cost = ((points - centers[0]) ** 2).sum(axis=1)
    for c in range(1, k):
        s = 0.0
        for j in range(m):
            s += float(cost[j]) * float(weights[j])
        r = rand.next_double() * s
        cum, j = 0.0, 0
        while j < m and cum < r:
            cum += float(weights[j]) * float(cost[j])
            j += 1
        centers[c] = points[0] if j == 0 else points[j - 1]
        cost = np.minimum(((points - centers[c]) ** 2).sum(axis=1), cost)
    old = np.full(m, -1)
    it, moved = 0, True
    while moved and it < max_iter:
        moved = False
        counts = np.zeros(k)
        sums = np.zeros((k, d))
        for i in range(m):
            idx = _find_closest(centers, points[i])
            sums[idx] += weights[i] * points[i]
            counts[idx] += weights[i]
            if idx != old[i]:
                moved = True
                old[i] = idx
        for j in range(k):
            if counts[j] == 0.0:
                centers[j] = points[rand.next_int(m)]
            else:
                centers[j] = sums[j] * (1.0 / counts[j])
        it += 1
    return centers


def xorshift_jump_table(max_log2: int = 42) -> np.ndarray:
    """Column form of T^(2^j), j = 0..max_log2-1, T = one XORShiftRandom.next step over GF(2).

    Entry [j, i] is T^(2^j) applied to the unit vector e_i, so T^(2^j) x = XOR of [j, i] over the
    set bits i of x.  The k-means|| sampling kernel uses it to jump a partition's stream to any row.
    """
    def step(x: int) -> int:
        x ^= (x << 21) & _M64
        x ^= x >> 35
        x ^= (x << 4) & _M64
        return x

    cols = [step(1 << i) for i in range(64)]
    out = np.zeros((max_log2, 64), dtype=np.uint64)

    def apply(cols_, x):
        y = 0
        i = 0
        while x:
            if x & 1:
                y ^= cols_[i]
            x >>= 1
            i += 1
        return y

    for j in range(max_log2):
        out[j] = np.array(cols, dtype=np.uint64)
        cols = [apply(cols, c) for c in cols]
    return out
