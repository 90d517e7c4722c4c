"""cProfile of one fit from given centres on a resident table: python tools/fit_profile.py n d k"""
import os, sys, cProfile, pstats, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hkm_b200
from hkm_b200 import datagen, _runtime as RT
from hkm_b200.ml.clustering import KMeans
n, d, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dev = RT.device()
X = datagen.hospital_matrix_device(n, d, min(k, 16), 20250331, dev)
C0 = X[:k].cpu().numpy()
est = KMeans(k=k, maxIter=20, tol=0.0, seed=42).setInitialCenters(C0)
m = est.fit({"features": X}); torch.cuda.synchronize()
t0 = time.perf_counter(); m = est.fit({"features": X}); torch.cuda.synchronize(); print("fit wall ms", (time.perf_counter() - t0) * 1e3, "numIter", m.summary.numIter)
pr = cProfile.Profile(); pr.enable()
m = est.fit({"features": X}); torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
