"""One default-init fit on a resident table (for an ncu launch list): python tools/fit_kernels.py n d k"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hkm_b200
from hkm_b200 import datagen, _runtime as RT
from hkm_b200.ml.clustering import KMeans
n, d, k = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dev = RT.device()
X = datagen.hospital_matrix_device(n, d, min(k, 16), 20250331, dev)
est = KMeans(k=k, maxIter=20, tol=0.0, seed=42)
m = est.fit({"features": X}); torch.cuda.synchronize()      # warm
torch.cuda.nvtx.range_push("fit"); m = est.fit({"features": X}); torch.cuda.synchronize(); torch.cuda.nvtx.range_pop()
print("numIter", m.summary.numIter, "cost", m.summary.trainingCost)
