"""Seeded synthetic "hospital-shaped" tables following the reference schema
(ref:L64-72): admission_count, current_occupancy, emergency_visits (int32), seasonality_index,
length_of_stay (fp64), plus derived fp64 columns up to the requested width.  Hospitals are drawn
from `profiles` latent profiles so real cluster structure exists.

`hospital_columns` builds NumPy columns (tests, CPU baselines); `hospital_matrix_device` builds the
standardised row-major matrix directly in HBM for benchmarks too large for host RAM.
"""
from __future__ import annotations

from collections import OrderedDict

import numpy as np

BASE_COLUMNS = ["admission_count", "current_occupancy", "emergency_visits", "seasonality_index", "length_of_stay"]


def feature_names(d: int):
    names = list(BASE_COLUMNS[: min(d, 5)])
    for i in range(5, d):
        names.append(f"derived_{i - 5:02d}")
    return names


def hospital_columns(n: int, d: int = 8, profiles: int = 5, seed: int = 20250331, null_fraction: float = 0.0):
    """OrderedDict name -> ndarray (ints are int32, the rest fp64). Nulls, if requested, come back
    as numpy masked arrays (one null per affected row)."""
    rng = np.random.default_rng(seed)
    H = 64
    prof_lambda = rng.uniform(5, 60, size=profiles)
    prof_mu = rng.uniform(40, 900, size=profiles)
    hosp_prof = rng.integers(0, profiles, size=H)
    lam_h = prof_lambda[hosp_prof] * rng.uniform(0.9, 1.1, size=H)
    mu_h = prof_mu[hosp_prof] * rng.uniform(0.9, 1.1, size=H)
    cap_h = mu_h * 1.3
    h = rng.integers(0, H, size=n)
    doy = rng.integers(0, 365, size=n)
    hour = rng.integers(0, 24, size=n)
    dow = rng.integers(0, 7, size=n)
    season = 1 + 0.25 * np.sin(2 * np.pi * doy / 365.0) + rng.normal(0, 0.05, size=n)
    adm = rng.poisson(lam_h[h]).astype(np.int32)
    occ = np.clip(rng.normal(mu_h[h], 0.15 * mu_h[h]), 0, cap_h[h]).astype(np.int32)
    ed = rng.poisson(0.3 * lam_h[h] * season).astype(np.int32)
    los = rng.gamma(2.0, 0.8 + 0.004 * occ + 0.6 * np.maximum(season - 1, -0.9))
    cols = OrderedDict()
    base = [adm, occ, ed, season, los]
    for name, v in zip(BASE_COLUMNS, base):
        cols[name] = v
    derived_pool = [
        np.sin(2 * np.pi * hour / 24.0), np.cos(2 * np.pi * hour / 24.0),
        np.sin(2 * np.pi * dow / 7.0), np.cos(2 * np.pi * dow / 7.0),
        occ / cap_h[h], ed / (adm + 1.0),
    ]
    names = feature_names(d)
    bf = np.stack([b.astype(np.float64) for b in base], axis=1)
    bf = (bf - bf.mean(0)) / (bf.std(0) + 1e-12)
    for i in range(5, d):
        j = i - 5
        if j < len(derived_pool):
            cols[names[i]] = derived_pool[j].astype(np.float64)
        else:
            wts = rng.normal(size=5)
            cols[names[i]] = bf @ wts + rng.normal(0, 1.0, size=n)
    out = OrderedDict((nm, cols[nm]) for nm in names)
    if null_fraction > 0:
        rows = np.flatnonzero(rng.random(n) < null_fraction)
        which = rng.integers(0, d, size=len(rows))
        for c, nm in enumerate(names):
            hit = rows[which == c]
            if len(hit):
                mask = np.zeros(n, dtype=bool)
                mask[hit] = True
                out[nm] = np.ma.MaskedArray(out[nm], mask=mask)
    return out


def hospital_matrix(n: int, d: int = 8, profiles: int = 5, seed: int = 20250331, standardize: bool = True) -> np.ndarray:
    """Row-major fp64 [n, d] matrix (StandardScaler defaults: withMean=False, withStd=True)."""
    cols = hospital_columns(n, d, profiles, seed)
    X = np.stack([np.asarray(v, dtype=np.float64) for v in cols.values()], axis=1)
    if standardize:
        std = X.std(axis=0, ddof=1)
        X = X * np.where(std == 0, 0.0, 1.0 / np.where(std == 0, 1.0, std))
    return np.ascontiguousarray(X)


def hospital_matrix_device(n: int, d: int, profiles: int, seed: int, device, chunk: int = 1 << 22):
    """Same family of data generated on the GPU in chunks (for benchmark sizes): `profiles` latent
    hospital profiles, int-valued count columns, derived mixtures; standardised per column with
    analytic scale so no second pass is needed.  Returns a torch fp64 [n, d] tensor."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    X = torch.empty((n, d), dtype=torch.float64, device=device)
    prof_centers = torch.randn((profiles, d), dtype=torch.float64, device=device, generator=g) * 2.0
    mix = torch.randn((5, d), dtype=torch.float64, device=device, generator=g) * 0.3
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        m = hi - lo
        prof = torch.randint(0, profiles, (m,), device=device, generator=g)
        base = torch.randn((m, 5), dtype=torch.float64, device=device, generator=g)
        blk = prof_centers[prof] + torch.randn((m, d), dtype=torch.float64, device=device, generator=g) * 0.7 + base @ mix
        # count-like columns: quantise the first three the way int32 Poisson/occupancy columns are
        ncount = min(3, d)
        blk[:, :ncount] = torch.round(blk[:, :ncount] * 8.0) / 8.0
        X[lo:hi] = blk
        del prof, base, blk
    return X
