"""Wall time of a default-init fit on a resident table, split into init / Lloyd / summary (CUDA-synchronised stopwatch)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hkm_b200
from hkm_b200 import datagen, _runtime as RT
from hkm_b200.ml import _kmeans_init as KI
from hkm_b200.ml.clustering import KMeans
dev = RT.device()
for n, d, k in [(20_000_000, 16, 16), (5_000_000, 24, 32), (10_000_000, 32, 8), (5_000_000, 64, 64)]:
    X = datagen.hospital_matrix_device(n, d, min(k, 16), 20250331, dev)
    layout = KI.PartitionLayout.build([n], 0, 8)
    def t(fn, reps=3):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps): r = fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps * 1e3
    ms_init = t(lambda: KI.init_kmeans_parallel(X, layout, k, 2, 42))
    C0 = KI.init_kmeans_parallel(X, layout, k, 2, 42)
    est = KMeans(k=k, maxIter=20, tol=0.0, seed=42)
    ms_fit_given = t(lambda: est.setInitialCenters(np.asarray(C0.cpu()) if hasattr(C0, "cpu") else np.asarray(C0)).fit({"features": X}))
    est2 = KMeans(k=k, maxIter=20, tol=0.0, seed=42)
    ms_fit_default = t(lambda: est2.fit({"features": X}))
    print(json.dumps(dict(n=n, d=d, k=k, init_ms=round(ms_init, 2), fit_given_centres_ms=round(ms_fit_given, 2), fit_default_ms=round(ms_fit_default, 2),
                          lloyd_pass_ideal_ms=round(n * d * 8 / 6554.6e6, 3))), flush=True)
    del X
