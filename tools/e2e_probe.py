import sys, time; sys.path.insert(0, "/root/repo")
import torch, numpy as np, hkm_b200
from hkm_b200 import datagen
from hkm_b200.ml.clustering import KMeans
dev = torch.device("cuda")
n, d, k = 10_000_000, 32, 8
X = datagen.hospital_matrix_device(n, d, k, 20250331, dev)
host = torch.empty((n, d), dtype=torch.float64, pin_memory=True); host.copy_(X)
init = X[:k].cpu().numpy()
km = KMeans(k=k, seed=42, maxIter=20, tol=0.0).setInitialCenters(init)
def t(fn, reps=3):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
print("h2d ms", t(lambda: host.to(dev, non_blocking=True)))
print("fit(device) ms", t(lambda: km.fit({"features": hkm_b200.DeviceMatrix(X)})))
print("fit(host) ms", t(lambda: km.fit({"features": host})))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable(); km.fit({"features": hkm_b200.DeviceMatrix(X)}); torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(14)
