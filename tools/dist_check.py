"""torchrun check of the row-sharded path on real GPUs (NCCL):  KMeans.fit + ClusteringEvaluator +
StandardScaler on shards must match the oracle run on the whole table (rank 0 checks).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/dist_check.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import hkm_b200
from hkm_b200 import datagen
from hkm_b200.ml.clustering import KMeans
from hkm_b200.ml.evaluation import ClusteringEvaluator
from hkm_b200.ml.feature import StandardScaler

ok = True
for n, d, k, seed in [(80_000, 8, 5, 42), (160_000, 24, 16, 7), (240_000, 20, 24, 3)]:
    P = 8
    raw = datagen.hospital_matrix(n, d=d, profiles=k, seed=20250331, standardize=False)
    lo, hi = rank * n // world, (rank + 1) * n // world                   # n divisible by P: shards line up with Spark's slices
    shard = {"raw": raw[lo:hi].copy()}
    sm = StandardScaler(inputCol="raw", outputCol="features").fit(shard)
    t = sm.transform(shard)
    model = KMeans(k=k, seed=seed, maxIter=20).setNumPartitions(P).fit(t)
    sil = ClusteringEvaluator().evaluate(model.summary.predictions)
    if rank == 0:
        from oracle import hkm_oracle as O
        mean, std = O.scaler_fit(raw, P=P)
        Xo = O.scaler_transform(raw, mean, std)
        ref = O.kmeans_fit(Xo, k, seed, P=P)
        e_std = np.abs(sm.std - std).max() / np.abs(std).max()
        e_c = np.abs(np.array(model.clusterCenters()) - ref["centers"]).max() / np.abs(ref["centers"]).max()
        e_cost = abs(model.summary.trainingCost - ref["training_cost"]) / ref["training_cost"]
        e_sil = abs(sil - O.silhouette(Xo, ref["labels"]))
        agree = (model.summary.predictions.to_numpy("prediction") == ref["labels"][lo:hi]).mean()
        good = e_std < 1e-11 and e_c < 1e-6 and e_cost < 1e-6 and e_sil < 1e-5 and agree >= 0.9999 and model.summary.numIter == ref["num_iter"] \
            and model.summary.clusterSizes == ref["cluster_sizes"].tolist()
        ok &= good
        print(f"world={world} n={n} d={d} k={k}: std_rel={e_std:.1e} centres_rel={e_c:.1e} cost_rel={e_cost:.1e} sil_abs={e_sil:.1e} "
              f"labels={agree:.6f} numIter={model.summary.numIter}/{ref['num_iter']} {'OK' if good else 'FAIL'}", flush=True)
dist.barrier()
if rank == 0:
    print("DIST ALL OK" if ok else "DIST FAILED", flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
