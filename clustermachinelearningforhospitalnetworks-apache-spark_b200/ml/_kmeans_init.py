"""KMeans initialisation: `random` (KMeans.initRandom) and `k-means||`
(KMeans.initKMeansParallel + LocalKMeans, 8a-6/8a-7), orchestrated on the host exactly as the
Spark driver does, with every pass over the rows in CUDA.

Both depend on Spark's partition layout by construction (per-partition RNG streams), so the
layout is explicit: each rank's local rows are cut into `P_local` ParallelCollectionRDD-style
slices and global partition index = (partitions of lower ranks) + local index.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List

import numpy as np
import torch

from .. import _native as N
from .. import _runtime as RT
from .. import _sparkrand as SR


@dataclass
class PartitionLayout:
    rank: int
    world: int
    rank_rows: List[int]          # rows per rank
    parts_per_rank: List[int]     # partitions per rank
    part_sizes: List[int]         # global list of partition sizes (all ranks, in order)

    @staticmethod
    def build(rank_rows: List[int], rank: int, num_partitions: int) -> "PartitionLayout":
        world = len(rank_rows)
        base = max(1, num_partitions // world)
        parts_per_rank = [base] * world
        for r in range(num_partitions - base * world):  # spread any remainder over the first ranks
            if num_partitions > base * world:
                parts_per_rank[r % world] += 1
        sizes: List[int] = []
        for r in range(world):
            sizes += SR.equal_partition_sizes(rank_rows[r], parts_per_rank[r])
        return PartitionLayout(rank, world, list(rank_rows), parts_per_rank, sizes)

    @property
    def n_global(self) -> int:
        return sum(self.rank_rows)

    @property
    def part_index0(self) -> int:
        return sum(self.parts_per_rank[: self.rank])

    @property
    def local_part_sizes(self) -> List[int]:
        i0 = self.part_index0
        return self.part_sizes[i0: i0 + self.parts_per_rank[self.rank]]

    @property
    def row_offset(self) -> int:
        return sum(self.rank_rows[: self.rank])


def _gather_global_rows(X: torch.Tensor, layout: PartitionLayout, global_idx: List[int]) -> np.ndarray:
    """Rows of the (rank-sharded) matrix by global index, in the given order, on every rank."""
    lo = layout.row_offset
    hi = lo + X.shape[0]
    mine = [(pos, g - lo) for pos, g in enumerate(global_idx) if lo <= g < hi]
    if mine:
        sel = torch.tensor([m[1] for m in mine], dtype=torch.int64, device=X.device)
        rows = X.index_select(0, sel).cpu().numpy()
    else:
        rows = np.zeros((0, X.shape[1]))
    parts = RT.all_gather_object(([m[0] for m in mine], rows))
    out = np.empty((len(global_idx), X.shape[1]))
    for pos, rws in parts:
        for p, r in zip(pos, rws):
            out[p] = r
    return out


def init_random(X: torch.Tensor, layout: PartitionLayout, k: int, seed: int) -> np.ndarray:
    """data.takeSample(false, k, new XORShiftRandom(seed).nextInt()).map(_.vector).distinct"""
    s2 = SR.XORShiftRandom(seed).next_int()
    idx = SR.take_sample(layout.part_sizes, k, s2)
    return SR.distinct_rows(_gather_global_rows(X, layout, idx))


def _unit(C: np.ndarray) -> np.ndarray:
    return C * (1.0 / np.sqrt((C * C).sum(axis=1)))[:, None]


def init_kmeans_parallel(X: torch.Tensor, layout: PartitionLayout, k: int, init_steps: int, seed: int,
                         first_idx: int = -1, debug: dict | None = None, X_unit: torch.Tensor | None = None) -> np.ndarray:
    """X_unit (cosine measure): the distance passes run on the unit rows against unit candidates -- there
    |x - c|^2 = 2 * cosine distance, and the sampling rule 2*k*cost/sum(cost) does not see the factor; the candidates
    themselves stay RAW rows and LocalKMeans stays Euclidean, as in Spark."""
    dev = X.device
    Xd = X if X_unit is None else X_unit
    prep = (lambda c: c) if X_unit is None else _unit
    n, d = X.shape
    seed2 = SR.XORShiftRandom(seed).next_int()
    if first_idx < 0:
        s = SR.take_sample(layout.part_sizes, 1, seed2)
        if not s:
            raise ValueError("No samples available from an empty dataset")
        first_idx = s[0]
    centers = _gather_global_rows(X, layout, [first_idx])
    new = centers.copy()

    costs = torch.empty(n, dtype=torch.float64, device=dev)
    N.fill_f64(costs, float("inf"))
    sum_costs = torch.zeros(1, dtype=torch.float64, device=dev)
    scratch = N.kmpar_new_scratch(d, dev)
    strip = N.kmpar_strip_rows()
    lsizes = layout.local_part_sizes
    part_rows = torch.tensor(SR.partition_bounds(lsizes), dtype=torch.int64, device=dev)
    strips = [(s + strip - 1) // strip for s in lsizes]
    strip_start = torch.tensor(SR.partition_bounds(strips), dtype=torch.int64, device=dev)
    total_strips = sum(strips)
    count = torch.zeros(1, dtype=torch.int32, device=dev)
    cap = max(4096, 64 * k)
    idx_buf = torch.empty(cap, dtype=torch.int64, device=dev)
    step_counts = []

    for step in range(init_steps):
        if len(new):
            N.kmpar_cost_update(Xd, torch.from_numpy(np.ascontiguousarray(prep(new))).to(dev), costs, sum_costs, scratch)
            RT.all_reduce_sum(sum_costs)
        # else: no new centres -> costs and their sum are unchanged (pointCost over nothing is +inf)
        while True:
            N.kmpar_sample(costs, part_rows, strip_start, total_strips, layout.part_index0, seed2, step, k, sum_costs, idx_buf, count)
            c = int(count.item())
            if c <= idx_buf.numel():
                break
            idx_buf = torch.empty(2 * c, dtype=torch.int64, device=dev)
        local = torch.sort(idx_buf[:c]).values
        rows = X.index_select(0, local).cpu().numpy() if c else np.zeros((0, d))
        gathered = RT.all_gather_object(rows)      # rank order == global partition order
        new = np.concatenate(gathered, axis=0) if gathered else rows
        step_counts.append(len(new))
        centers = np.concatenate([centers, new], axis=0)

    cand = SR.distinct_rows(centers)
    if debug is not None:
        debug.update(candidates=cand.copy(), step_counts=step_counts, first_idx=first_idx)
    if len(cand) <= k:
        return cand
    m = len(cand)
    labels_work = torch.empty(n, dtype=torch.int32, device=dev)
    mind2_work = torch.empty(n, dtype=torch.float64, device=dev)
    counts = torch.zeros(m, dtype=torch.int64, device=dev)
    N.closest_counts(Xd, torch.from_numpy(np.ascontiguousarray(prep(cand))).to(dev), labels_work, mind2_work, counts)
    RT.all_reduce_sum(counts)
    weights = counts.cpu().numpy().astype(np.float64)
    if debug is not None:
        debug.update(weights=weights.copy())
    return N.local_kmeanspp(0, cand, weights, k, 30)
