"""Small-shape workload for compute-sanitizer (memcheck / racecheck / synccheck): every kernel family of libhkm once.

  compute-sanitizer --tool memcheck  python tools/sanitize.py
  compute-sanitizer --tool racecheck python tools/sanitize.py
"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
N = importlib.import_module("clustermachinelearningforhospitalnetworks-apache-spark_b200._native")
import hkm_b200
from hkm_b200.ml.clustering import KMeans
from hkm_b200.ml.evaluation import ClusteringEvaluator
from hkm_b200.ml.feature import StandardScaler, VectorAssembler

rng = np.random.default_rng(0)
only = sys.argv[1] if len(sys.argv) > 1 else "all"


def one_pass(n, d, k, variant, w=False, mind2=False):
    X = torch.from_numpy(rng.normal(size=(n, d)) + rng.integers(0, k, size=(n, 1)) * 2.0).cuda()
    C = X[:k].clone()
    partial = torch.empty(k * d + k + 1, dtype=torch.float64, device="cuda")
    labels = torch.empty(n, dtype=torch.int32, device="cuda")
    m2 = torch.empty(n, dtype=torch.float64, device="cuda") if mind2 else None
    wt = torch.from_numpy(rng.uniform(0.5, 2.0, size=n)).cuda() if w else None
    N.set_lloyd_variant(variant)
    try:
        N.assign_reduce(X, C, partial, labels, m2, weights=wt, scratch=N.new_scratch(d, k, "cuda"))
        N.assign_reduce(X, C, None, labels, do_reduce=False)
    finally:
        N.set_lloyd_variant(0)
    torch.cuda.synchronize()
    print(f"pass n={n} d={d} k={k} variant={variant} weights={w} mind2={mind2}: counts sum {int(partial[k * d:k * d + k].sum().item())}", flush=True)


if only in ("all", "tp"):
    # pipelined tcgen05 variant: DMMA reduce, private sets, owned shared set, every warp budget
    for n, d, k, v in [(700, 16, 16, 4), (700, 24, 12, 4), (700, 64, 16, 50), (700, 24, 32, 40), (700, 32, 64, 40), (700, 64, 64, 42), (700, 48, 24, 42),
                       (333, 7, 33, 40), (700, 16, 16, 54), (700, 64, 64, 53), (700, 20, 50, 52)]:
        one_pass(n, d, k, v)
if only in ("all", "fp64"):
    # fp64 tensor-pipe kernel, gather kernel (plain / weighted / distances), warp-specialised and the round-1 tcgen05 variant
    one_pass(900, 32, 8, 5); one_pass(900, 8, 5, 5); one_pass(900, 16, 16, 5); one_pass(900, 24, 32, 5); one_pass(900, 64, 40, 5, mind2=True)
    one_pass(900, 12, 20, 1); one_pass(900, 9, 40, 0, w=True); one_pass(900, 24, 32, 0, mind2=True); one_pass(900, 96, 5, 0)
if only in ("all", "general"):
    one_pass(600, 128, 8, 0); one_pass(600, 20, 100, 0); one_pass(300, 200, 70, 0, w=True)
if only in ("all", "fit"):
    # estimator: loader, scaler, k-means|| (cost update, sampling, candidate weights), Lloyd loop, summary, silhouette
    cols = hkm_b200.datagen.hospital_columns(3000, d=8, profiles=5, seed=1, null_fraction=0.01)
    t = VectorAssembler(inputCols=list(cols), outputCol="raw", handleInvalid="skip").transform(dict(cols))
    t = StandardScaler(inputCol="raw", outputCol="features").fit(t).transform(t)
    for k in (5, 20):
        m = KMeans(k=k, seed=42, maxIter=5).fit(t)
        s = ClusteringEvaluator().evaluate(m.transform(t))
        print(f"fit k={k}: cost {m.summary.trainingCost:.3f} silhouette {s:.4f}", flush=True)
    m = KMeans(k=4, seed=1, maxIter=3, distanceMeasure="cosine").fit(t)
    print("cosine fit", m.summary.trainingCost, flush=True)
print("sanitize workload done", flush=True)
