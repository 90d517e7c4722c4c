"""cfg5 (BASELINE configs[4]): model-selection sweep k=2..32 on one resident 5M x 24 table -- where does the time go?
Prints per-k wall times (ms, after a synchronize) of init / Lloyd / summary pass / silhouette."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hkm_b200
from hkm_b200 import datagen
from hkm_b200.ml import clustering as CL, _kmeans_init as KI, evaluation as EV
from hkm_b200 import _runtime as RT

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
d, kmax = 24, 32
dev = RT.device()
X = datagen.hospital_matrix_device(n, d, 8, 20250331, dev)
layout = KI.PartitionLayout.build([n], 0, 8)

def tick():
    torch.cuda.synchronize(); return time.perf_counter()

tot = dict(init=0.0, lloyd=0.0, summary=0.0, sil=0.0)
for rep in range(2):
    tot = dict(init=0.0, lloyd=0.0, summary=0.0, sil=0.0)
    rows = []
    for k in range(2, kmax + 1):
        t0 = tick(); C0 = KI.init_kmeans_parallel(X, layout, k, 2, 42)
        t1 = tick(); C, it, cost = CL.lloyd(X, C0, None, 20, 1e-4)
        t2 = tick(); lab, sizes = CL._predict_and_count(X, torch.from_numpy(C).to(dev))
        t3 = tick(); s = EV.silhouette(X, lab)
        t4 = tick()
        rows.append((k, it, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3, s))
        for key, v in zip(tot, (t1 - t0, t2 - t1, t3 - t2, t4 - t3)): tot[key] += v * 1e3
for r in rows: print("k=%2d iters=%2d init=%7.2f lloyd=%7.2f summary=%6.2f sil=%6.2f silhouette=%.4f" % r)
print(json.dumps({"n": n, "d": d, "sweep_ms": {k: round(v, 1) for k, v in tot.items()}, "total_ms": round(sum(tot.values()), 1)}))
