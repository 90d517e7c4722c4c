"""Where k-means|| init time goes (5M x 24, k=32)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hkm_b200
from hkm_b200 import datagen, _native as N, _sparkrand as SR, _runtime as RT
from hkm_b200.ml import _kmeans_init as KI
n, d = 5_000_000, 24
k = int(sys.argv[1]) if len(sys.argv) > 1 else 32
dev = RT.device()
X = datagen.hospital_matrix_device(n, d, 8, 20250331, dev)
layout = KI.PartitionLayout.build([n], 0, 8)
import cProfile, pstats
KI.init_kmeans_parallel(X, layout, k, 2, 42)
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(5): KI.init_kmeans_parallel(X, layout, k, 2, 42)
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
